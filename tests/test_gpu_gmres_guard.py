"""The accuracy guard of the GMRES driver (iemic_b200/csrc/gmres_guard.h) on the GPU, through the C ABI.

THCM without fixed pressure points has a singular Jacobian (the constant-pressure null vector).  At a converged state of the
3x6x6 2-D MOC box the residual F is rounding noise with a component outside the range of J; GMRES in exact arithmetic
stagnates at that component.  The reference's modified Gram-Schmidt (GMRESSolver.H:177-181) stagnates too; the CGS2 driver
keeps an orthogonal basis, its recurrence residual falls below the tolerance while the explicit residual is O(1) and the
solution is a large multiple of the null vector.  With the guard the solve returns 2 and an iterate close to the
minimum-norm least-squares solution."""
import numpy as np
import pytest
import scipy.sparse as sp

import moc2d
import oracle_lib as ol
from test_known_answer_2dmoc import OracleBackend

pytestmark = pytest.mark.gpu


def _converged_state():
    cfg, mask, B = moc2d.setup(OracleBackend)
    rp, col = B.graph()
    x = moc2d.continuation(B, np.zeros(rp.size - 1), rp, col, "COMB", 0.5, 0.1)
    x, ok = moc2d.newton(B, x, rp, col, tol=1e-12)
    assert ok
    return cfg, mask, B, x


def test_guard_returns_a_bounded_iterate_on_the_singular_jacobian():
    import torch
    import iemic_b200 as ib
    cfg, mask, B, x = _converged_state()
    G = ib.Ocean(cfg, mask)
    import iemic_b200.grids as grids
    for k, v in grids.STARTING_PARS["2dmoc"].items():
        G.setPar(k, v)
    G.setPar("COMB", 0.5)
    F = G.computeRHS_host(x)
    val, _ = G.computeJacobian_host(x)
    rp, col = G.graph()
    n = rp.size - 1
    J = sp.csr_matrix((val, col, rp), shape=(n, n)).toarray()
    assert np.linalg.svd(J, compute_uv=False)[-1] < 1e-10 * np.abs(J).max()   # singular
    nb = np.linalg.norm(F)
    assert 0 < nb < 1e-10
    xl = np.linalg.lstsq(J, F, rcond=1e-11)[0]
    d_binv = torch.zeros(n * 6, dtype=torch.float64, device="cuda")
    G.block_jacobi(d_binv)
    d_b = torch.from_numpy(F).cuda()
    for ortho in (ib.CGS2, ib.CGS2F):
        d_x = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        rc, hist, it = G.solve(d_b, d_x, m=500, maxit=4000, tol=1e-10, ortho=ortho, prec=ib.PREC_BLOCKJACOBI, d_blockinv=d_binv)
        xs = d_x[:n].cpu().numpy()
        assert rc == 2, (rc, it)
        # unguarded, the "converged" answer has norm O(0.1 .. 10); the least-squares solution is of the order of ||F||
        assert np.linalg.norm(xs) <= 1e3 * max(np.linalg.norm(xl), nb), (np.linalg.norm(xs), np.linalg.norm(xl))
        assert np.linalg.norm(F - J @ xs) <= 1.0001 * nb
    # a right-hand side in the range of J is solved to the tolerance as before
    b = J @ np.ones(n)
    d_b = torch.from_numpy(b).cuda()
    d_x = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    rc, hist, it = G.solve(d_b, d_x, m=500, maxit=4000, tol=1e-10, ortho=ib.CGS2, prec=ib.PREC_BLOCKJACOBI, d_blockinv=d_binv)
    assert rc == 0
    xs = d_x[:n].cpu().numpy()
    assert np.linalg.norm(b - J @ xs) <= 2e-10 * np.linalg.norm(b)
