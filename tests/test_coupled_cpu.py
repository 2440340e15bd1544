"""Row N2 on the CPU: the ocean side of the coupling to an atmosphere / sea-ice model (Coupled Temperature = 1).

The product's host code (parameters, per-cell surface forcing, coefficient tables) and its __host__ __device__ row code --
the arithmetic the CUDA kernels run -- against the oracle's restatement of usrc.F90:237-333, 725-741, forcing.F90:73-96 and
inserts.F90: bit-exact, like every other case of tests/test_host_rows.py."""
import numpy as np
import pytest

import cases
import hostcheck_lib as hl
import oracle_lib as ol


def _pair(name):
    cfg, mask, pars = cases.case(name)
    O, H = ol.Oracle(cfg, mask), hl.HostCheck(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        H.set_par(k, v)
    return cfg, mask, O, H


def _same(O, H, cfg, mask):
    O.rhs(np.zeros(O.ndim))  # the oracle masks Frc during an evaluation
    assert np.array_equal(O.frc(), H.frc())
    for kind, seed in (("random", 1), ("smooth", 2)):
        x = cases.state(cfg, mask, seed=seed, kind=kind)
        assert np.array_equal(O.rhs(x), H.rhs(x))
        vo, bo = O.jac(x)
        vh, bh = H.jac(x)
        assert O.dropped() == 0
        assert np.array_equal(vo, vh)
        assert np.array_equal(bo, bh)


@pytest.mark.parametrize("name", ["coupled_small", "coupled_box"])
def test_attached_atmosphere_and_sea_ice_bit_exact(name):
    cfg, mask, O, H = _pair(name)
    _same(O, H, cfg, mask)                                  # nothing attached yet: only the solar term of the forcing
    x = cases.state(cfg, mask, seed=1)
    F0, (v0, _) = O.rhs(x), O.jac(x)
    cases.attach(O, cfg, seaice=False)
    cases.attach(H, cfg, seaice=False)
    assert O.getdeps() == H.getdeps() and O.getdeps()["lvsc"] != 0.0
    _same(O, H, cfg, mask)                                  # atmosphere: latent heat term dedt, per-cell forcing
    F1, (v1, _) = O.rhs(x), O.jac(x)
    assert not np.array_equal(F0, F1) and not np.array_equal(v0, v1)
    cases.attach(O, cfg, seaice=True)
    cases.attach(H, cfg, seaice=True)
    _same(O, H, cfg, mask)                                  # sea ice: mask-dependent centre coefficient and the (T,S) entry
    F2, (v2, _) = O.rhs(x), O.jac(x)
    assert not np.array_equal(F1, F2) and not np.array_equal(v1, v2)
    # the T row of an ice-covered surface cell now depends on S (Al(5,TT,SS) = -QTnd*zeta*a0*mc)
    rp, col = O.graph()
    rows = np.repeat(np.arange(rp.size - 1), np.diff(rp))
    ts = (rows % 6 == 4) & (col == rows + 1)
    assert np.count_nonzero(v2[ts] != v1[ts]) > 0
    # a parameter change afterwards keeps lvsc / nus as the last synchronisation left them (usrc.F90:163-181 calls forcing and lin only)
    O.set_par("COMB", 0.35)
    H.set_par("COMB", 0.35)
    assert O.getdeps() == H.getdeps()
    _same(O, H, cfg, mask)


def test_insert_takes_effect_at_the_next_parameter_call():
    """inserts.F90 only stores the field; forcing and lin run in set_*_parameters / setparcs (Ocean::synchronize, Ocean.C:1443-1493)"""
    cfg, mask, O, H = _pair("coupled_small")
    f = cases.coupling_fields(cfg)
    before = H.frc().copy()
    for obj in (O, H):
        obj.insert("tatm", f["tatm"])
        obj.insert("msi", f["msi"])
    assert np.array_equal(H.frc(), before)
    x = cases.state(cfg, mask, seed=4)
    assert np.array_equal(O.rhs(x), H.rhs(x))               # the sea-ice mask is not in the coefficients yet either
    for obj in (O, H):
        obj.set_par("TEMP", 9.0)
    assert not np.array_equal(H.frc(), before)
    _same(O, H, cfg, mask)


def test_uncoupled_model_ignores_the_atmosphere_fields():
    cfg, mask, O, H = _pair("natl8")
    f = cases.coupling_fields(cfg)
    x = cases.state(cfg, mask, seed=5)
    F0 = H.rhs(x)
    for obj in (O, H):
        obj.insert("tatm", f["tatm"])                       # coupled_T = 0: inserts.F90:24 ignores it
        obj.insert("albe", f["albe"])
        obj.set_par("TEMP", obj.get_par("TEMP"))
    assert np.array_equal(H.rhs(x), F0)
    assert np.array_equal(O.rhs(x), F0)


def test_coupled_salinity_fields_are_refused():
    cfg, mask, pars = cases.case("coupled_small")
    H = hl.HostCheck(cfg, mask)
    f = cases.coupling_fields(cfg)
    H.insert("patm", f["qatm"])                             # coupled_S = 0: ignored, as in inserts.F90:91
    with pytest.raises(RuntimeError):
        hl.HostCheck(dict(cfg, coupled_S=1), mask)


@pytest.mark.parametrize("name", ["coupled_small", "coupled_box"])
def test_derivative_field_and_coupling_blocks(name):
    """m_probe::get_derivatives (probe.F90:429-444: the expression the CUDA kernel runs, compiled for the CPU) and
    Ocean::getBlock(Atmosphere / SeaIce) (Ocean.C:1538-1746) against the oracle / a restatement on the oracle's numbers"""
    cfg, mask, O, H = _pair(name)
    f = cases.attach(O, cfg)
    cases.attach(H, cfg)
    x = cases.state(cfg, mask, seed=5)
    assert np.array_equal(O.suno(), H.suno())
    d = O.dftdm(x)
    assert np.array_equal(d, H.dftdm(x))
    surf = np.asarray(mask).reshape(cfg["l"] + 2, cfg["m"] + 2, cfg["n"] + 2)[cfg["l"], 1:-1, 1:-1].ravel()
    assert np.all(d[surf != 0] == 0.0) and np.count_nonzero(d[surf == 0]) == np.count_nonzero(surf == 0)
    Ao, So = cases.blocks_from_oracle(O, cfg, mask, x, f, cases.ATMOS_PARS["da"])
    assert cases.compact(H.atmos_block()) == Ao
    assert cases.compact(H.seaice_block(d)) == So
    assert len(Ao) == np.count_nonzero(surf == 0)
    # an ice-covered point is cut off from the atmosphere (1 - M = 0), open water feels all three atmosphere unknowns
    full = [e for (row, e), M in zip(Ao, f["msi"][surf == 0]) if M == 1.0]
    open_ = [e for (row, e), M in zip(Ao, f["msi"][surf == 0]) if M == 0.0]
    assert full and all(all(a == 0.0 for _, a in e) for e in full)
    assert open_ and all(all(a != 0.0 for _, a in e) for e in open_)


def test_get_derivatives_refreshes_nus_like_the_fortran():
    """probe.F90:414 assigns the module variable nus from the current parameters; set_par alone does not (usrc.F90:163-181)"""
    cfg, mask, O, H = _pair("coupled_small")
    cases.attach(O, cfg)
    cases.attach(H, cfg)
    nus0 = H.getdeps()["nus"]
    for obj in (O, H):
        obj.set_par("COMB", 0.6)
    assert H.getdeps()["nus"] == nus0 == O.getdeps()["nus"]
    x = cases.state(cfg, mask, seed=5)
    assert np.array_equal(O.dftdm(x), H.dftdm(x))
    assert O.getdeps() == H.getdeps() and H.getdeps()["nus"] != nus0


def test_coupling_block_known_answer_of_the_reference_test():
    """src/tests/test_coupled.C:402-447: C12 applied to a constant atmosphere vector gives, in the T row of an ocean surface point
    without sea ice, -(Ooa + lvsc*eta*qdim - comb*sunp*So*da) * value within 1e-7."""
    cfg, mask, O, H = _pair("coupled_box")
    cases.attach(O, cfg, seaice=False)
    cases.attach(H, cfg, seaice=False)
    n, m, l = cfg["n"], cfg["m"], cfg["l"]
    d = H.getdeps()
    rows, beg, jco, co = H.atmos_block()
    suno = H.suno()
    surf = np.asarray(mask).reshape(l + 2, m + 2, n + 2)[l, 1:-1, 1:-1]
    checked = 0
    for value in (1.234, 2.12, -23.5):
        v = np.full(3 * n * m, value)
        for r in range(rows.size):
            cell = rows[r] // 6
            i, j = cell % n, (cell // n) % m
            assert rows[r] % 6 == 4 and cell // (n * m) == l - 1 and surf[j, i] == 0
            got = sum(co[p] * v[jco[p]] for p in range(beg[r], beg[r + 1]))
            want = -(d["Ooa"] + d["lvsc"] * d["eta"] * d["qdim"] - H.get_par("COMB") * H.get_par("SUNP") * suno[j] * cases.ATMOS_PARS["da"]) * value
            assert abs(got - want) <= 1e-7 * max(1.0, abs(want))
            checked += 1
    assert checked == 3 * np.count_nonzero(surf == 0)


def test_coupling_blocks_are_the_derivatives_of_the_residual():
    """The blocks against the oracle's F by differences: -d F / d(atmosphere unknown) and -dF/dM at a handful of surface points
    (the sign convention of Ocean::getBlock: "negating as the Jacobian is taken negative" means block = d F_THCM / d unknown)."""
    cfg, mask, O, H = _pair("coupled_small")
    f = cases.attach(O, cfg)
    cases.attach(H, cfg)
    n, m, l = cfg["n"], cfg["m"], cfg["l"]
    x = cases.state(cfg, mask, seed=5)
    A = dict((r, dict(e)) for r, e in cases.compact(H.atmos_block()))
    S = dict((r, dict(e)) for r, e in cases.compact(H.seaice_block(H.dftdm(x))))

    def F_with(field, values, seaice):
        O.insert(field, values)
        (O.set_seaice_parameters if seaice else O.set_atmos_parameters)(cases.SEAICE_PARS if seaice else cases.ATMOS_PARS)
        return O.rhs(x)

    rng = np.random.default_rng(0)
    ocean_rows = sorted(A)
    for row in rng.choice(ocean_rows, size=6, replace=False):
        sr = (row // 6) % (n * m)
        for field, colv, blk, seaice, h in (("tatm", 3 * sr + 0, A, False, 1e-3), ("qatm", 3 * sr + 1, A, False, 1e-7), ("albe", 3 * sr + 2, A, False, 1e-4),
                                              ("msi", 4 * sr + 2, S, True, 1e-5)):
            base = f[field].copy()
            up, dn = base.copy(), base.copy()
            up[sr] += h
            dn[sr] -= h
            dF = (F_with(field, up, seaice) - F_with(field, dn, seaice)) / (2 * h)
            F_with(field, base, seaice)
            want = blk[row][colv]
            # the sea-ice mask also sits in the operator (mc terms of lin): F = A(M) x - Frc(M); the block holds the forcing part
            # plus the operator part evaluated at this state, which is what dftdm = QTos - QToa contains (probe.F90:433-442)
            assert abs(dF[row] - want) <= 1e-6 * max(abs(want), 1.0), (field, dF[row], want)
            others = np.delete(dF, row)
            if field != "msi":
                assert np.abs(others).max() <= 1e-6 * max(abs(want), 1.0)
