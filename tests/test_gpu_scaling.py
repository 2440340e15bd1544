"""THCM block scaling and the point-block-Jacobi builder on the GPU, through the C ABI, against the oracle's
restatement of m_scaling / THCM::RecomputeScaling (pinned against LAPACK in tests/test_scaling_cpu.py).

Tolerances: the averaged 6x6 block is a sum over all ocean cells taken in another order than the Fortran loop: 1e-12 of
its largest entry (plus what the tanh entries of the T,S rows carry, 1e-6 of the row scale / number of cells); the
scaling factors follow from a 6x6 inverse: 1e-9 relative (1e-6 for the T,S factors); LeftScale / Multiply / Reciprocal are one rounding each:
bit-identical to numpy."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

CASES = ["natl8", "natl8_sres1", "test6x6x4", "2dmoc", "2dmoc_run", "rand_periodic", "rand_box", "rand_fixp", "global4"]


def _diag_blocks(rp, col, val, nrows):
    nb = nrows // 6
    blocks = np.zeros((nb, 6, 6))
    rows = np.repeat(np.arange(nrows), np.diff(rp))
    sel = (col // 6) == (rows // 6)
    blocks[rows[sel] // 6, rows[sel] % 6, col[sel] % 6] = val[sel]
    return blocks


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("kind", ["random", "smooth"])
def test_average_block_and_scaling(name, kind):
    import torch
    cfg, mask, O, G = cases.make_pair(name)
    x = cases.state(cfg, mask, seed=4, kind=kind)
    O.jac(x)
    G.computeJacobian_host(x, want_values=False)
    db_o = O.average_block()
    db_g = G.average_block()
    scale = np.abs(db_o).max()
    plain = np.ones((6, 6), bool)
    plain[4:, 4:] = False  # T,S x T,S: forward differences of tanh
    assert np.abs(db_o - db_g)[plain].max() <= 1e-12 * scale
    assert np.abs(db_o - db_g).max() <= 1e-6 * scale
    ref = O.recompute_scaling()
    d_r = torch.zeros(G.nrows, dtype=torch.float64, device="cuda")
    d_c = torch.zeros(G.nrows, dtype=torch.float64, device="cuda")
    if ref is None:
        import iemic_b200 as ib
        with pytest.raises(ib.B200Error):
            G.recompute_scaling(d_r, d_c)
        return
    rs, cs = G.recompute_scaling(d_r, d_c)
    G.sync()
    ts = np.arange(G.nrows) % 6 >= 4  # factors that depend on the tanh finite differences of the averaged block
    for got, exp in ((d_r.cpu().numpy(), ref[0]), (d_c.cpu().numpy(), ref[1])):
        np.testing.assert_allclose(got[~ts], exp[~ts], rtol=1e-9)
        np.testing.assert_allclose(got[ts], exp[ts], rtol=1e-6)
    assert rs[4] == rs[5] and cs[4] == cs[5]


@pytest.mark.parametrize("name", ["natl8", "rand_periodic", "global4"])
def test_scale_problem_roundtrip(name):
    """Ocean::scaleProblem / unscaleProblem (Ocean.C:1163-1250): LeftScale by rowScaling, rhs .*= rowScaling, and back"""
    import torch
    cfg, mask, O, G = cases.make_pair(name)
    x = cases.state(cfg, mask, seed=5)
    v0, _ = G.computeJacobian_host(x)
    rp, col = G.graph()
    d_r = torch.zeros(G.nrows, dtype=torch.float64, device="cuda")
    G.recompute_scaling(d_r, None)
    r = d_r.cpu().numpy()
    rows = np.repeat(np.arange(G.nrows), np.diff(rp))
    G.left_scale(d_r)
    G.sync()
    v1, _ = G.values_host()
    assert np.array_equal(v1, v0 * r[rows])
    rhs = torch.rand(G.nrows, dtype=torch.float64, device="cuda")
    rhs0 = rhs.clone()
    G.multiply(1.0, d_r, rhs, 0.0, rhs)
    d_ri = torch.empty_like(d_r)
    G.reciprocal(d_r, d_ri)
    G.sync()
    assert torch.equal(rhs, d_r * rhs0)
    assert torch.equal(d_ri, 1.0 / d_r)
    # the scaled operator applied to a vector equals the scaled result of the unscaled one
    import oracle_lib as ol
    w = np.random.default_rng(3).uniform(-1, 1, G.nrows)
    y1 = G.applyMatrix_host(w)
    assert np.array_equal(y1, ol.spmv(rp, col, v1, w))
    G.left_scale(d_ri)
    G.sync()
    v2, _ = G.values_host()
    np.testing.assert_allclose(v2, v0, rtol=4e-16, atol=0)
    acc = rhs.clone()
    G.multiply(2.0, d_ri, rhs, -1.0, acc)  # acc = 2*(1/r).*rhs - acc
    G.sync()
    assert torch.allclose(acc, 2.0 * (d_ri * rhs) - rhs, rtol=0, atol=1e-15)


@pytest.mark.parametrize("name", CASES)
def test_block_jacobi_inverses(name):
    import torch
    cfg, mask, O, G = cases.make_pair(name)
    x = cases.state(cfg, mask, seed=6)
    vg, _ = G.computeJacobian_host(x)
    rp, col = G.graph()
    blocks = _diag_blocks(rp, col, vg, G.nrows)
    d_b = torch.zeros(G.nrows * 6, dtype=torch.float64, device="cuda")
    nsing = G.block_jacobi(d_b)
    binv = d_b.cpu().numpy().reshape(-1, 6, 6)
    # blocks with an exactly zero pivot column come back as the identity and are counted
    sing = np.all(binv == np.eye(6), axis=(1, 2)) & ~np.all(blocks == np.eye(6), axis=(1, 2))
    assert nsing == int(sing.sum())
    cond = np.array([np.linalg.cond(b) for b in blocks])
    ok = ~sing & (cond < 1e12)
    assert ok.sum() >= 0.9 * (~sing).sum()
    prod = np.einsum("bij,bjk->bik", binv[ok], blocks[ok])
    err = np.abs(prod - np.eye(6)).max(axis=(1, 2))
    assert np.all(err <= 1e-13 * np.maximum(cond[ok], 1.0)), (err / cond[ok]).max()


@pytest.mark.parametrize("name", ["natl8_sres1", "rand_periodic", "global4"])
def test_gmres_with_device_built_preconditioner(name):
    """same preconditioner blocks on both sides: the residual history must match the oracle's GMRESSolver restatement"""
    import torch
    import oracle_lib as ol
    import iemic_b200 as ib
    cfg, mask, O, G = cases.make_pair(name)
    x = cases.state(cfg, mask, seed=6)
    vg, _ = G.computeJacobian_host(x)
    rp, col = G.graph()
    b = ol.spmv(rp, col, vg, np.ones(G.nrows))
    d_binv = torch.zeros(G.nrows * 6, dtype=torch.float64, device="cuda")
    G.block_jacobi(d_binv)
    binv = d_binv.cpu().numpy()
    m = 20
    rc_o, xo, ho, it_o = ol.gmres(rp, col, vg, b, np.zeros(G.nrows), m, m - 1, 0.0, blockinv=binv)
    d_b = torch.from_numpy(b).cuda()
    d_x = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    rc_g, hg, it_g = G.solve(d_b, d_x, m=m, maxit=m - 1, tol=0.0, ortho=ib.MGS, prec=ib.PREC_BLOCKJACOBI, d_blockinv=d_binv, check_every=0)
    assert it_o == it_g
    ho, hg = ho[:m + 1], hg[:m + 1]
    assert np.all(np.abs(ho - hg) <= 1e-10 * np.maximum(np.abs(ho), 1e-300) + 1e-14), (ho, hg)
