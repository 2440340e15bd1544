"""THCM block scaling on the CPU: the oracle's restatement of m_scaling (src/ocean/scaling.F90) and
THCM::RecomputeScaling (src/ocean/THCM.C:1694-1750), pinned against LAPACK itself (the dgetrf / dgetri the reference
calls, reached through scipy), and the product's host code for the same path (thcm::scal, Host::diag_info, the side
buffer of the rows the C++ glue overwrites) against the oracle."""
import numpy as np
import pytest
from scipy.linalg import lapack

import cases
import hostcheck_lib as hl
import oracle_lib as ol

CASES = ["natl8", "natl8_sres1", "test6x6x4", "2dmoc", "rand_periodic", "rand_box", "rand_fixp", "global_half"]


def _scal_lapack(db):
    """scaling.F90:198-284 with the real LAPACK routines"""
    mat = np.asfortranarray(np.array(db, np.float64))
    lu, piv, info = lapack.dgetrf(mat)
    assert info == 0
    mat, info = lapack.dgetri(lu, piv)
    assert info == 0
    m = lambda i, j: mat[i - 1, j - 1]
    dr, dc = np.zeros(6), np.zeros(6)
    dr[0] = dc[0] = 1.0
    idc = np.sqrt(m(1, 1) / m(2, 2))
    dr[1] = dc[1] = 1 / idc
    idr = np.sqrt(abs(m(1, 1) / m(4, 4)))
    dr[3] = dc[3] = 1 / idr
    if abs(m(4, 3)) > abs(m(3, 3)):
        idr = m(1, 1) / (idr * m(4, 3))
    else:
        idr = np.sqrt(abs(m(1, 1) / m(3, 3)))
    dr[2] = dc[2] = 2 / idr
    if abs(m(4, 5) * m(5, 4)) < .01 * abs(m(4, 4) * m(5, 5)):
        mat[3, 4] = 1
        mat[4, 3] = 1
    idc = np.sqrt(abs(m(1, 1) * m(4, 5) / (m(5, 4) * m(5, 5))))
    idr = m(1, 1) / (idc * m(5, 5))
    dr[4], dc[4] = 1 / idr, 1 / idc
    if abs(m(4, 6) * m(6, 4)) < .01 * abs(m(4, 4) * m(6, 6)):
        mat[3, 5] = 1
        mat[5, 3] = 1
    idc = np.sqrt(abs(m(1, 1) * m(4, 6) / (m(6, 4) * m(6, 6))))
    idr = m(1, 1) / (idc * m(6, 6))
    dr[5], dc[5] = 1 / idr, 1 / idc
    return dr, dc


def _pair(name):
    cfg, mask, pars = cases.case(name)
    O, H = ol.Oracle(cfg, mask), hl.HostCheck(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        H.set_par(k, v)
    return cfg, mask, O, H


def _average_block_numpy(O, x):
    """scaling.F90:29-62 once more, in numpy, from the Fortran-level CSR"""
    beg, jco, co, _ = O.fortran_matrix(x)
    n, m, l = O.cfg["n"], O.cfg["m"], O.cfg["l"]
    lm = O.landm.reshape(l + 2, m + 2, n + 2)[1:-1, 1:-1, 1:-1].ravel()  # interior cells, i fastest
    rows = np.repeat(np.arange(O.ndim), np.diff(beg))
    cols = jco - 1
    sel = (rows // 6 == cols // 6) & (lm[rows // 6] == 0)
    db = np.zeros((6, 6))
    np.add.at(db, (rows[sel] % 6, cols[sel] % 6), co[sel])
    nl = int((lm == 0).sum())
    return db / nl if nl else db


@pytest.mark.parametrize("seed", range(6))
def test_scal_matches_lapack(seed):
    rng = np.random.default_rng(seed)
    db = rng.standard_normal((6, 6)) + 3.0 * np.eye(6)
    db[0, 0], db[1, 1] = 2.0 + rng.random(), 2.0 + rng.random()  # mat(1,1)/mat(2,2) > 0 like the friction blocks
    ref = _scal_lapack(db)
    if not np.all(np.isfinite(np.concatenate(ref))):
        pytest.skip("sqrt of a negative ratio: undefined in the reference too")
    for got in (ol.scal(db), hl.scal(db)):
        assert got is not None
        np.testing.assert_allclose(got[0], ref[0], rtol=1e-11)
        np.testing.assert_allclose(got[1], ref[1], rtol=1e-11)


def test_scal_singular_block():
    db = np.zeros((6, 6))
    db[0, 0] = 1.0
    assert ol.scal(db) is None and hl.scal(db) is None


@pytest.mark.parametrize("name", CASES)
def test_average_block_and_scaling(name):
    cfg, mask, O, H = _pair(name)
    for kind, seed in (("random", 1), ("smooth", 2)):
        x = cases.state(cfg, mask, seed=seed, kind=kind)
        db_np = _average_block_numpy(O, x)
        db_o = O.average_block()
        scale = np.abs(db_np).max()
        np.testing.assert_allclose(db_o, db_np, rtol=0, atol=1e-13 * scale)
        got = H.scaling(x)
        ref = O.recompute_scaling()
        if ref is None:
            assert got is None
            continue
        db_h, rs, cs = got
        np.testing.assert_allclose(db_h, db_o, rtol=0, atol=1e-13 * scale)
        # LAPACK on the oracle's block
        dr, dc = _scal_lapack(db_o)
        r6, c6 = 1.0 / dr, 1.0 / dc
        r6[4] = r6[5] = 0.5 * (r6[4] + r6[5])
        c6[4] = c6[5] = 0.5 * (c6[4] + c6[5])
        np.testing.assert_allclose(rs, r6, rtol=1e-9)
        np.testing.assert_allclose(cs, c6, rtol=1e-9)
        # the full vectors: ocean rows carry the six factors, every other row 1
        rowS, colS = ref
        n, m, l = cfg["n"], cfg["m"], cfg["l"]
        ocean = np.asarray(mask).reshape(l + 2, m + 2, n + 2)[1:-1, 1:-1, 1:-1].ravel() == 0
        exp_r = np.where(np.repeat(ocean, 6), np.tile(rs, ocean.size), 1.0)
        exp_c = np.where(np.repeat(ocean, 6), np.tile(cs, ocean.size), 1.0)
        np.testing.assert_allclose(rowS, exp_r, rtol=1e-9)
        np.testing.assert_allclose(colS, exp_c, rtol=1e-9)
