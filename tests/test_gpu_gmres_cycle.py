"""The one-launch GMRES cycles (iemic_b200/csrc/gmres_cluster.cu, gmres_cycle.cu) against the kernel-per-operation driver.

All three run the algorithm of GMRESSolver.H:142-213 with the same orthogonalisation; they differ in the order of the sums, so
the recurrence residuals agree to rounding while the solve is well conditioned.  IEMIC_GMRES_CYCLE is read on every solve:
"0" = one kernel per operation, "coop" = cooperative kernel only, "cluster" = cluster kernel only, unset = automatic."""
import os

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _solve(G, b, binv, mode, ortho, m, maxit, tol, check_every):
    import torch
    old = os.environ.get("IEMIC_GMRES_CYCLE")
    if mode is None:
        os.environ.pop("IEMIC_GMRES_CYCLE", None)
    else:
        os.environ["IEMIC_GMRES_CYCLE"] = mode
    try:
        d_x = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        import iemic_b200 as ib
        l0 = G.launch_count()
        rc, hist, it = G.solve(b, d_x, m=m, maxit=maxit, tol=tol, ortho=ortho, prec=ib.PREC_BLOCKJACOBI if binv is not None else ib.PREC_NONE,
                               d_blockinv=binv, check_every=check_every)
        G.sync()
        return rc, hist[:it + 2].copy(), it, d_x[:G.nrows].cpu().numpy(), G.launch_count() - l0
    finally:
        if old is None:
            os.environ.pop("IEMIC_GMRES_CYCLE", None)
        else:
            os.environ["IEMIC_GMRES_CYCLE"] = old


def _problem(name, prec):
    import torch
    import iemic_b200 as ib
    cfg, mask, pars = cases.case(name)
    G = ib.Ocean(cfg, mask)
    for k, v in pars.items():
        G.setPar(k, v)
    x = cases.state(cfg, mask, seed=6)
    G.computeJacobian(torch.from_numpy(x).cuda())
    ones = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    ones[:G.nrows] = 1.0
    b = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    G.applyMatrix(ones, b)
    binv = None
    if prec:
        binv = torch.zeros(G.nrows * 6, dtype=torch.float64, device="cuda")
        G.block_jacobi(binv)
    return G, b, binv


# natl8 (1536 rows, dense integral row) fits one cluster; rand_periodic (17094 rows) is the cooperative kernel's size
@pytest.mark.parametrize("name,modes", [("natl8", (None, "cluster", "coop")), ("natl8_sres1", (None, "coop")), ("rand_periodic", (None, "coop"))])
@pytest.mark.parametrize("ortho", ["cgs2", "cgs2f"])
@pytest.mark.parametrize("prec", [0, 1])
def test_one_launch_cycle_matches_the_per_kernel_driver(name, modes, ortho, prec):
    import iemic_b200 as ib
    G, b, binv = _problem(name, prec)
    o = {"cgs2": ib.CGS2, "cgs2f": ib.CGS2F}[ortho]
    m, maxit = 25, 74 # three full cycles, no convergence test
    rc0, h0, it0, x0, l0 = _solve(G, b, binv, "0", o, m, maxit, 0.0, 0)
    for mode in modes:
        if mode == "coop" and ortho != "cgs2f":
            continue # the cooperative kernel implements the fused variant only
        rc1, h1, it1, x1, l1 = _solve(G, b, binv, mode, o, m, maxit, 0.0, 0)
        assert (rc1, it1) == (rc0, it0)
        assert np.all(np.abs(h1 - h0) <= 1e-8 * np.abs(h0) + 1e-13), (mode, h0, h1)
        assert np.abs(x1 - x0).max() <= 1e-7 * max(np.abs(x0).max(), 1.0)
        if mode is not None or name != "rand_periodic" or ortho == "cgs2f":
            assert l1 < l0 / 3, (mode, l1, l0) # one launch per cycle instead of several per iteration


@pytest.mark.parametrize("name", ["natl8", "rand_periodic"])
def test_one_launch_cycle_stops_where_the_reference_stops(name):
    """convergence inside a cycle: same iteration count, return code and solution as the driver that tests every iteration"""
    import iemic_b200 as ib
    G, b, binv = _problem(name, 1)
    _, h, _, _, _ = _solve(G, b, binv, "0", ib.CGS2F, 40, 39, 0.0, 0)
    # a tolerance between two consecutive residuals of the first cycle, where the history drops fastest
    j = 10 + int(np.argmax(h[9:34] / h[10:35]))
    assert h[j] < (1.0 - 1e-5) * h[j - 1] # the two paths agree to 1e-8: the gap leaves no room for a different stopping point
    tol = float(0.5 * (h[j] + h[j - 1]))
    rc0, h0, it0, x0, _ = _solve(G, b, binv, "0", ib.CGS2F, 40, 400, tol, 1)
    rc1, h1, it1, x1, _ = _solve(G, b, binv, None, ib.CGS2F, 40, 400, tol, 8)
    assert rc0 == 0 and rc1 == 0
    assert it0 == it1 == j # GMRESSolver.H leaves its loop before counting the converged iteration
    assert np.all(np.abs(h1 - h0) <= 1e-8 * np.abs(h0) + 1e-13)
    assert np.abs(x1 - x0).max() <= 1e-7 * max(np.abs(x0).max(), 1.0)


def test_maxit_inside_a_cycle_and_mgs_keeps_the_per_kernel_path():
    import iemic_b200 as ib
    G, b, binv = _problem("natl8", 0)
    rc0, h0, it0, x0, _ = _solve(G, b, binv, "0", ib.CGS2, 30, 41, 0.0, 8)
    rc1, h1, it1, x1, _ = _solve(G, b, binv, None, ib.CGS2, 30, 41, 0.0, 8)
    assert (rc0, it0) == (rc1, it1) and it1 == 42
    assert np.all(np.abs(h1 - h0) <= 1e-8 * np.abs(h0) + 1e-13)
    # modified Gram-Schmidt is the reference's dependent chain: it has no one-launch variant
    rc2, h2, it2, x2, l2 = _solve(G, b, binv, None, ib.MGS, 30, 41, 0.0, 8)
    assert it2 == 42 and l2 > 200
