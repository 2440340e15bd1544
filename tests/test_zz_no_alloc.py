"""vec_b200_dot_noalloc (include/iemic_b200.h): the dot product whose reduction scratch is a static device array -- no
allocation in the call.  Checked through the C ABI against the existing entry point vec_b200_dot on two calls in a row
(the scratch keeps the previous call's contents).  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def test_dot_noalloc_matches_vec_dot_twice():
    import torch
    import iemic_b200 as ib
    L = ib.lib()
    cfg, mask, O, G = cases.make_pair("global_half")
    n = G.nrows
    rng = np.random.default_rng(7)
    out = torch.full((4,), -123.0, dtype=torch.float64, device="cuda")
    for call in range(2):
        a = torch.from_numpy(rng.uniform(-1, 1, n)).cuda()
        b = torch.from_numpy(rng.uniform(-1, 1, n)).cuda()
        ref = G.dot(a, b)                                     # vec_b200_dot: handle-owned (allocated) scratch
        torch.cuda.synchronize()
        rc = L.vec_b200_dot_noalloc(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_int64(n), C.c_void_p(out.data_ptr() + 8 * call), None)
        assert rc == 0, L.iemic_b200_last_error()
        torch.cuda.synchronize()
        got = float(out[call].item())
        exact = float(np.dot(a.cpu().numpy(), b.cpu().numpy()))
        assert abs(got - ref) <= 1e-12 * max(abs(ref), 1.0), (call, got, ref)
        assert abs(got - exact) <= 1e-10 * max(abs(exact), 1.0)
    assert float(out[2].item()) == -123.0                      # nothing written beyond d_out[0] of each call
    # odd length and unaligned vectors take the scalar tail
    a = torch.from_numpy(rng.uniform(-1, 1, 1001)).cuda()
    rc = L.vec_b200_dot_noalloc(C.c_void_p(a.data_ptr() + 8), C.c_void_p(a.data_ptr() + 8), C.c_int64(999), C.c_void_p(out.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert abs(float(out[0].item()) - float((a[1:1000] ** 2).sum().item())) <= 1e-12 * 1000


def test_dot_noalloc_rejects_oversize():
    import torch
    import iemic_b200 as ib
    L = ib.lib()
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = L.vec_b200_dot_noalloc(C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()), C.c_int64((1 << 30) + 1), C.c_void_p(x.data_ptr()), None)
    assert rc < 0 and b"outside" in L.iemic_b200_last_error()
    rc = L.vec_b200_dot_noalloc(None, C.c_void_p(x.data_ptr()), C.c_int64(8), C.c_void_p(x.data_ptr()), None)
    assert rc < 0
