"""Domain decomposition on the CPU: every rank's host-side setup (thcm_host.cpp: Decomp2D blocks, halo ring storage,
local graph, forcing, signatures) and the row arithmetic evaluated on its local storage must reproduce the single-domain
oracle exactly -- the `||J*1||` rank-count invariance idiom of src/tests/test_matrix.C:148-192, sharpened to bit level.
Also runs a real world_size-2 gloo exchange using the halo plan."""
import os
import subprocess
import sys

import numpy as np
import pytest
import scipy.sparse as sp

import cases
import hostcheck_lib as hl
import oracle_lib as ol


def _assemble(name, nranks):
    attach = name.endswith("+coupling")   # row N2: an atmosphere and sea ice attached (every rank gets the whole interface fields)
    name = name.replace("+coupling", "")
    cfg, mask, pars = cases.case(name)
    O = ol.Oracle(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
    if attach:
        cases.attach(O, cfg)
    x = cases.state(cfg, mask, seed=7)
    N = O.ndim
    F = np.full(N, np.nan)
    B = np.full(N, np.nan)
    rows_g, cols_g, vals_g = [], [], []
    seen = np.zeros(N, int)
    for r in range(nranks):
        H = hl.HostCheck(cfg, mask, rank=r, nranks=nranks)
        for k, v in pars.items():
            H.set_par(k, v)
        if attach:
            cases.attach(H, cfg)
        gid = H.gid()
        assert np.all(gid[:H.ndim] >= 0)
        seen[gid[:H.ndim]] += 1
        xloc = np.where(gid >= 0, x[np.maximum(gid, 0)], 0.0)
        Fl = H.rhs(xloc)
        F[gid[:H.ndim]] = Fl
        val, diagB = H.jac(xloc)
        B[gid[:H.ndim]] = diagB
        rp, col = H.graph()
        lr = np.repeat(np.arange(H.ndim), np.diff(rp))
        assert np.all(gid[col] >= 0), "a stored column points at an unused halo slot"
        rows_g.append(gid[lr]); cols_g.append(gid[col]); vals_g.append(val)
        frc = H.frc()
        O.rhs(x)
        assert np.array_equal(frc, O.frc()[gid[:H.ndim]])
    assert np.all(seen == 1), "every global row is owned by exactly one rank"
    rows_g, cols_g, vals_g = np.concatenate(rows_g), np.concatenate(cols_g), np.concatenate(vals_g)
    return cfg, O, x, F, B, rows_g, cols_g, vals_g


@pytest.mark.parametrize("name,nranks", [("rand_periodic", 2), ("rand_periodic", 4), ("rand_periodic", 8), ("rand_box", 2), ("rand_box", 6),
                                         ("global_half", 8), ("coupled4", 4), ("coupled_small+coupling", 4), ("coupled_box+coupling", 6), ("natl8_sres1", 4), ("2dmoc_sres1", 2), ("2dmoc_run_sres1", 4)])
def test_rank_pieces_reproduce_the_global_problem(name, nranks):
    cfg, O, x, F, B, rows_g, cols_g, vals_g = _assemble(name, nranks)
    assert np.array_equal(F, O.rhs(x))
    vo, bo = O.jac(x)
    assert np.array_equal(B, bo)
    rp, col = O.graph()
    ro = np.repeat(np.arange(O.ndim), np.diff(rp))
    # same pattern and values: sort both by (row, col)
    ko = np.lexsort((col, ro))
    kg = np.lexsort((cols_g, rows_g))
    assert rows_g.size == ro.size
    assert np.array_equal(rows_g[kg], ro[ko]) and np.array_equal(cols_g[kg], col[ko])
    assert np.array_equal(vals_g[kg], vo[ko])
    # the test_matrix.C idiom
    J = sp.csr_matrix((vals_g, (rows_g, cols_g)), shape=(O.ndim, O.ndim))
    Jo = sp.csr_matrix((vo, col, rp), shape=(O.ndim, O.ndim))
    assert abs(np.linalg.norm(J @ np.ones(O.ndim)) - np.linalg.norm(Jo @ np.ones(O.ndim))) < 1e-8


def test_decomposition_matches_reference_rule():
    from iemic_b200 import grids
    cfg, mask, pars = cases.case("global_half")
    for nranks in (1, 2, 4, 8):
        npN, npM = grids.decomp2d(cfg["n"], cfg["m"], nranks)
        cells = 0
        for r in range(nranks):
            H = hl.HostCheck(cfg, mask, rank=r, nranks=nranks)
            i0, j0, nloc, mloc, pn, pm = H.block()
            assert (pn, pm) == (npN, npM)
            assert i0 == (r % npN) * (cfg["n"] // npN) + min(cfg["n"] % npN, r % npN)
            cells += nloc * mloc
        assert cells == cfg["n"] * cfg["m"]


def test_decomposition_literal_table():
    """Decomp2D results worked out BY HAND from the reference's loop (src/trios/TRIOS_Domain.C:88-108: scan t1 = nprocs..1,
    t2 = nprocs/t1, keep the LAST pair with t1*t2 == nprocs and |m/t1 - n/t2| <= r_min, r_min starting at 100), as npN x npM."""
    from iemic_b200 import grids
    table = {
        (360, 160): {1: (1, 1), 2: (2, 1), 4: (4, 1), 6: (3, 2), 8: (4, 2)},
        (96, 38): {1: (1, 1), 2: (2, 1), 4: (4, 1), 8: (4, 2)},
        (48, 19): {2: (2, 1), 4: (4, 1), 8: (4, 2)},
    }
    for (n, m), rows in table.items():
        for nranks, want in rows.items():
            assert grids.decomp2d(n, m, nranks) == want, (n, m, nranks)
    # the product's C++ side (thcm::Host) must give the same blocks
    cfg, mask, pars = cases.case("global_half")
    for nranks, (npN, npM) in table[(48, 19)].items():
        H = hl.HostCheck(cfg, mask, rank=nranks - 1, nranks=nranks)
        assert H.block()[4:6] == (npN, npM)
    # offsets and sizes of the last block: remainders go to the first ranks (TRIOS_Domain.C:137-153); 48 = 4 x 12, 19 = 10 + 9
    H = hl.HostCheck(cfg, mask, rank=7, nranks=8)
    assert H.block()[:4] == (36, 10, 12, 9)


def test_sres0_multirank_dense_row_is_split_by_owner():
    """SRES = 0 on several ranks (THCM.C:2494-2513, the reference's partest_matrix idiom): the rank that owns the integral row keeps
    it in its pattern with the S unknowns it owns; the other ranks' share reaches it through one all-reduced double (capi.cu
    spmv_impl / rhs_epilogue), so no rank stores a column it does not hold."""
    cfg, mask, pars = cases.case("natl8")
    O = ol.Oracle(cfg, mask)
    rpo, colo = O.graph()
    dense = int(np.argmax(np.diff(rpo)))
    assert dense == O.rowintcon()
    for nranks in (2, 4, 8):
        owners = 0
        for r in range(nranks):
            H = hl.HostCheck(cfg, mask, rank=r, nranks=nranks)
            gid = H.gid()
            rp, col = H.graph()
            assert np.all(gid[col] >= 0)
            lens = np.diff(rp)
            rows = gid[:H.ndim]
            if dense in rows:
                owners += 1
                lr = int(np.flatnonzero(rows == dense)[0])
                assert H.rowintcon() == dense                                                # the GLOBAL row index, on every rank
                cols = gid[col[rp[lr]:rp[lr + 1]]]
                assert np.array_equal(np.sort(cols), np.sort(rows[rows % 6 == 5]))          # exactly the S unknowns this rank owns
                assert set(cols.tolist()) <= set(colo[rpo[dense]:rpo[dense + 1]].tolist())
            else:
                assert H.rowintcon() == dense and lens.max() <= 24
        assert owners == 1


def test_gloo_world2_halo_exchange():
    """Two real processes (gloo): exchange halos using the plan, then assemble; compared with the oracle inside the worker."""
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29541")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(here, "dist_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout
