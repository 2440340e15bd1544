"""Known answer I7 of the reference (src/tests/reft_ocean.C:50-80, test/ocean/ocean_reference.h5): the golden state the
reference itself wrote at the end of its 16^3 gateway continuation (Combined Forcing -> 0.02, Newton tolerance 1e-2,
P_VC = 9.95: convective adjustment active, land mask, periodic, forcing type 2).  A converged state of the reference
must be a root of the reference's F -- so the residual of the oracle, of the product's row code and of the CUDA path at
that state, with the parameters stored beside it, has to vanish to the reference's own convergence level.  Measured:
||F||_2 = 7.7e-6 against ||F(0)||_2 = 19.8 (momentum rows 1e-11, continuity 1e-15)."""
import numpy as np
import pytest

import cases


def _check_root(F, F0):
    assert np.linalg.norm(F0) > 10.0                      # the problem is forced: x = 0 is far from a root
    assert np.linalg.norm(F) <= 1.0e-5                    # Newton tolerance of the run that wrote the state: 1e-2
    assert np.abs(F[0::6]).max() <= 1e-9 and np.abs(F[1::6]).max() <= 1e-9   # u, v
    assert np.abs(F[2::6]).max() <= 1e-8                  # w
    assert np.abs(F[3::6]).max() <= 1e-13                 # continuity
    assert np.abs(F[4::6]).max() <= 1e-5 and np.abs(F[5::6]).max() <= 1e-6   # T, S


def test_oracle_and_row_code_vanish_at_the_reference_state():
    import hostcheck_lib as hl
    import oracle_lib as ol
    cfg, mask, pars, x = cases.reference_state()
    O, H = ol.Oracle(cfg, mask), hl.HostCheck(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        H.set_par(k, v)
    # per-field norms of the state, the quantities reft_ocean.C compares (u, v, T, S to 1e-3)
    assert abs(np.linalg.norm(x[4::6]) - 3.5162057673711735) < 1e-9
    F, F0 = O.rhs(x), O.rhs(np.zeros_like(x))
    _check_root(F, F0)
    Fh = H.rhs(x)
    ts = np.arange(F.size) % 6 >= 4
    assert np.array_equal(F[~ts], Fh[~ts])
    assert np.abs(F - Fh).max() <= 1e-12 * np.abs(F0).max()
    # the Jacobian at the root: Newton-consistent with F (one forward difference along a random direction over the
    # unknowns the reference's solution leaves free; the pinned ones -- land, no-slip corners, among them u,v(N,j,k) beside
    # land at i=1, which usol does not zero (usrc.F90:1031-1032) although `boundaries` pins them -- are exactly 0 in it)
    rp, col = O.graph()
    val, _ = O.jac(x)
    d = np.random.default_rng(0).uniform(-1, 1, x.size) * (x != 0.0)
    eps = 1e-8
    Jd = ol.spmv(rp, col, val, d)
    fd = (O.rhs(x + eps * d) - F) / eps
    assert np.abs(Jd - fd).max() <= 1e-4 * np.abs(Jd).max()


@pytest.mark.gpu
def test_cuda_path_vanishes_at_the_reference_state():
    import iemic_b200 as ib
    import oracle_lib as ol
    cfg, mask, pars, x = cases.reference_state()
    O, G = ol.Oracle(cfg, mask), ib.Ocean(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        G.setPar(k, v)
    F, F0 = G.computeRHS_host(x), G.computeRHS_host(np.zeros_like(x))
    _check_root(F, F0)
    Fo = O.rhs(x)
    ts = np.arange(F.size) % 6 >= 4
    assert np.array_equal(Fo[~ts], F[~ts])
    assert np.abs(Fo - F).max() <= 1e-12 * np.abs(F0).max()
