"""Preconditioner hook of gmres_b200_solve (gmres_b200_set_precond, the counterpart of Ocean.C:972-975 / GMRESSolver.H:153-172),
the documented length of the history array, and the global NormInf."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _blockinv(G):
    import torch
    d = torch.empty(G.nrows * 6, dtype=torch.float64, device="cuda")
    G.block_jacobi(d)
    return d


@pytest.mark.parametrize("ortho", ["mgs", "cgs2f"])
def test_callback_reproduces_block_jacobi_bit_for_bit(ortho):
    import torch
    import iemic_b200 as ib
    L = ib.lib()
    cfg, mask, O, G = cases.make_pair("global_half")
    x = cases.state(cfg, mask, seed=6)
    G.computeJacobian_host(x)
    binv = _blockinv(G)
    ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
    b = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    G.applyMatrix(ones, b)
    om = {"mgs": ib.MGS, "cgs2f": ib.CGS2F}[ortho]
    m = 20
    x1 = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    rc1, h1, it1 = G.solve(b, x1, m=m, maxit=2 * m - 1, tol=0.0, ortho=om, prec=ib.PREC_BLOCKJACOBI, d_blockinv=binv, check_every=0)
    calls = []

    def apply(v, z, n, stream):
        calls.append(n)
        return L.vec_b200_block_jacobi_apply(G.h, C.c_void_p(binv.data_ptr()), C.c_void_p(v), C.c_void_p(z))

    G.set_precond(apply)
    x2 = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    rc2, h2, it2 = G.solve(b, x2, m=m, maxit=2 * m - 1, tol=0.0, ortho=om, prec=ib.capi.PREC_CALLBACK, check_every=0)
    assert len(calls) == it2 and all(n == G.nrows for n in calls)
    assert (rc1, it1) == (rc2, it2)
    assert np.array_equal(h1, h2)
    assert torch.equal(x1[:G.nrows], x2[:G.nrows])
    # a failing callback ends the solve with an error, not a crash
    G.set_precond(lambda v, z, n, st: 7)
    with pytest.raises(ib.B200Error, match="callback returned 7"):
        G.solve(b, x2, m=m, maxit=m - 1, tol=0.0, ortho=om, prec=ib.capi.PREC_CALLBACK, check_every=0)
    G.set_precond(None)
    with pytest.raises(ib.B200Error, match="without gmres_b200_set_precond"):
        G.solve(b, x2, m=m, maxit=m - 1, tol=0.0, ortho=om, prec=ib.capi.PREC_CALLBACK, check_every=0)


def test_history_array_of_the_documented_length():
    """include/iemic_b200.h: hist holds maxit + 2 doubles.  A non-converged solve fills exactly that and nothing behind it."""
    import torch
    import iemic_b200 as ib
    L = ib.lib()
    cfg, mask, O, G = cases.make_pair("global_half")
    x = cases.state(cfg, mask, seed=6)
    G.computeJacobian_host(x)
    ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
    b = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    G.applyMatrix(ones, b)
    for m, maxit in ((10, 24), (7, 6), (5, 0)):
        hist = np.full(maxit + 2 + 4, -777.0)
        it = C.c_int(0)
        xs = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        rc = L.gmres_b200_solve(G.h, C.c_void_p(b.data_ptr()), C.c_void_p(xs.data_ptr()), m, maxit, 0.0, ib.CGS2F, ib.PREC_NONE, None, 1,
                                hist.ctypes.data_as(C.POINTER(C.c_double)), C.byref(it))
        assert rc >= 0 and it.value == maxit + 1
        assert np.all(hist[:maxit + 2] != -777.0) and np.all(hist[maxit + 2:] == -777.0)


def test_norminf_and_stream_ordered_copies():
    import torch
    import iemic_b200 as ib
    L = ib.lib()
    cfg, mask, O, G = cases.make_pair("natl8")
    a = np.random.default_rng(3).uniform(-1, 1, G.nrows)
    a[G.nrows // 3] = -5.25
    d = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    ib.capi.check(L.thcm_b200_upload(G.h, C.c_void_p(d.data_ptr()), a.ctypes.data_as(C.c_void_p), G.nrows))
    r = C.c_double(0.0)
    ib.capi.check(L.vec_b200_norminf(G.h, C.c_void_p(d.data_ptr()), C.byref(r)))
    assert r.value == 5.25
    back = np.zeros(G.nrows)
    ib.capi.check(L.thcm_b200_download(G.h, back.ctypes.data_as(C.c_void_p), C.c_void_p(d.data_ptr()), G.nrows))
    assert np.array_equal(back, a)
