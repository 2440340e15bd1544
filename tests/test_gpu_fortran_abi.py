"""Seam S1: the Fortran-named entry points (include/thcm_fortran_abi.h) against the oracle's Fortran-level outputs --
B of rhs_, and begA/jcoA/coA/coB of matrix_ in fillcolA's own format (1-based, |a| > 1e-10, slot-major order)."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["natl8", "natl8_sres1", "2dmoc", "rand_periodic", "rand_box", "global4"])
def test_fortran_symbols_match_oracle(name):
    import oracle_lib as ol
    import iemic_b200 as ib
    from fortran_abi import FortranTHCM
    cfg, mask, pars = cases.case(name)
    O = ol.Oracle(cfg, mask)
    T = FortranTHCM(cfg, mask)
    try:
        for k, v in pars.items():
            O.set_par(k, v)
            T.setpar(ib.PAR[k], v)
            assert T.getpar(ib.PAR[k]) == v
        x = cases.state(cfg, mask, seed=12)
        Bo, Bt = O.fortran_rhs(x), T.rhs(x)
        ts = (np.arange(Bo.size) % 6) >= 4
        assert np.array_equal(Bo[~ts], Bt[~ts])
        assert np.abs(Bo - Bt).max() <= 1e-12 * np.abs(Bo).max()
        bo, jo, co, cbo = O.fortran_matrix(x)
        bt, jt, ct, cbt = T.matrix(x)
        assert np.array_equal(cbo, cbt)
        rows = np.repeat(np.arange(bo.size - 1), np.diff(bo))
        tracer = (rows % 6 >= 4) & ((jo - 1) % 6 >= 4)
        if np.array_equal(bo, bt):
            assert np.array_equal(jo, jt), "same entries in the same (slot-major) order"
            assert np.array_equal(co[~tracer], ct[~tracer])
            scale = np.maximum.reduceat(np.abs(co), (bo[:-1] - 1)[np.diff(bo) > 0])
            full = np.zeros(bo.size - 1); full[np.diff(bo) > 0] = scale
            assert (np.abs(co - ct) / np.maximum(full[rows], 1e-300)).max() <= 1e-6
        else:
            # a finite-difference mixing entry on one side of the 1e-10 filter: compare as matrices
            import scipy.sparse as sp
            n = bo.size - 1
            Ao = sp.csr_matrix((co, jo - 1, bo - 1), shape=(n, n)); At = sp.csr_matrix((ct, jt - 1, bt - 1), shape=(n, n))
            assert abs(Ao - At).max() <= 1e-6 * abs(Ao).max()
        # integral-condition weights (thcm_utils.F90:285-323) against the oracle's coefficient vector
        val, ind = T.intcond_scaling()
        c = O.intcond_coeff()
        assert np.array_equal(c[ind - 1], val) and np.count_nonzero(c) == val.size
        # m_scaling::average_block / compute as THCM::RecomputeScaling drives them (THCM.C:1707-1745)
        db_o, ref = O.average_block(), O.recompute_scaling()
        if ref is not None:
            db_t, r_t, c_t = T.recompute_scaling()
            assert np.abs(db_o - db_t).max() <= 1e-6 * np.abs(db_o).max()
            np.testing.assert_allclose(r_t, ref[0], rtol=1e-6)
            np.testing.assert_allclose(c_t, ref[1], rtol=1e-6)
    finally:
        T.close()


def test_setsres_switches_surface_condition():
    """usrc.F90:417-429: setsres_ re-evaluates forcing + lin with the other salinity condition, parameters kept."""
    import oracle_lib as ol
    import iemic_b200 as ib
    from fortran_abi import FortranTHCM
    cfg, mask, pars = cases.case("natl8_sres1")
    T = FortranTHCM(cfg, mask)
    try:
        for k, v in pars.items():
            T.setpar(ib.PAR[k], v)
        T.setsres(0)
        cfg0 = dict(cfg, sres=0)
        O = ol.Oracle(cfg0, mask)
        for k, v in pars.items():
            O.set_par(k, v)
            assert T.getpar(ib.PAR[k]) == v
        x = cases.state(cfg, mask, seed=13)
        Bo, Bt = O.fortran_rhs(x), T.rhs(x)
        assert np.abs(Bo - Bt).max() <= 1e-12 * np.abs(Bo).max()
    finally:
        T.close()


def test_coupling_symbols_match_oracle():
    """row N2 through the Fortran names: m_inserts, set_atmos_parameters_ / set_seaice_parameters_, getdeps_, m_probe::get_derivatives"""
    import oracle_lib as ol
    import iemic_b200 as ib
    from fortran_abi import FortranTHCM
    cfg, mask, pars = cases.case("coupled_small_nomix")
    O = ol.Oracle(cfg, mask)
    T = FortranTHCM(cfg, mask)
    try:
        for k, v in pars.items():
            O.set_par(k, v)
            T.setpar(ib.PAR[k], v)
        cases.attach(O, cfg)
        cases.attach(T, cfg)
        assert O.getdeps() == T.getdeps()
        x = cases.state(cfg, mask, seed=12)
        assert np.array_equal(O.fortran_rhs(x), T.rhs(x))          # Mixing = 0: no tanh, every bit
        bo, jo, co, cbo = O.fortran_matrix(x)
        bt, jt, ct, cbt = T.matrix(x)
        assert np.array_equal(bo, bt) and np.array_equal(jo, jt) and np.array_equal(co, ct) and np.array_equal(cbo, cbt)
        dftdm, dfsdq, dfsdm, dfsdg = T.get_derivatives(x)
        assert np.array_equal(dftdm, O.dftdm(x))
        assert not dfsdq.any() and not dfsdm.any() and not dfsdg.any()   # Coupled Salinity = 0 (probe.F90:446)
    finally:
        T.close()

