"""Row N2 on the GPU, through the C ABI: the ocean side of the coupling to an atmosphere / sea-ice model (Coupled Temperature = 1).

Inserted interface fields and the models' parameters (Ocean::synchronize, Ocean.C:1443-1493) change the surface T forcing per cell
(forcing.F90:73-96) and the surface T row of the operator (usrc.F90:725-741).  Bars as in test_gpu_parity.py: bit-identical wherever
no tanh() is involved -- everything, in the Mixing = 0 case -- and 1e-12 / 1e-6 of the row scale in rows with the convective-
adjustment term.  Also: dF_T/dM (probe.F90:429-444), Ocean::getBlock(Atmosphere / SeaIce) (Ocean.C:1538-1746) against a restatement
on the oracle's numbers, CouplingBlock::applyMatrix and the accumulation of CoupledModel::applyMatrix."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _check(O, G, cfg, mask, strict):
    import torch
    O.rhs(np.zeros(O.ndim))  # the oracle masks Frc during an evaluation
    assert np.array_equal(O.frc(), G.forcing_host())
    rp, col = O.graph()
    rows = np.repeat(np.arange(rp.size - 1), np.diff(rp))
    for kind, seed in (("random", 3), ("smooth", 4)):
        x = cases.state(cfg, mask, seed=seed, kind=kind)
        Fo, (vo, bo) = O.rhs(x), O.jac(x)
        assert O.dropped() == 0
        Fg = G.computeRHS_host(x)
        vg, bg = G.computeJacobian_host(x)
        assert np.array_equal(bo, bg)
        if strict:
            assert np.array_equal(Fo, Fg) and np.array_equal(vo, vg)
        else:
            ts = np.arange(Fo.size) % 6 >= 4
            assert np.array_equal(Fo[~ts], Fg[~ts])
            assert np.abs(Fo - Fg).max() <= 1e-12 * np.abs(Fo).max()
            tracer = (rows % 6 >= 4) & (col % 6 >= 4)
            assert np.array_equal(vo[~tracer], vg[~tracer])
            full = np.zeros(rp.size - 1)
            full[np.diff(rp) > 0] = np.maximum.reduceat(np.abs(vo), rp[:-1][np.diff(rp) > 0])
            assert (np.abs(vo - vg) / np.maximum(full[rows], 1e-300)).max() <= 1e-6
        # the fused kernel (Jacobian + residual in one pass) gives the bits of the two stand-alone kernels
        dx = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        dx[:G.nrows] = torch.from_numpy(x).cuda()
        dF = torch.full((G.nvec,), 555.0, dtype=torch.float64, device="cuda")
        G.assemble(dx, dF, jacobian=True)
        G.sync()
        vf, bf = G.values_host()
        assert np.array_equal(dF[:G.nrows].cpu().numpy(), Fg) and np.array_equal(vf, vg) and np.array_equal(bf, bg)
    return Fo, vo


@pytest.mark.parametrize("name,strict", [("coupled_small_nomix", True), ("coupled_small", False), ("coupled_box", False)])
def test_attached_atmosphere_and_sea_ice(name, strict):
    cfg, mask, O, G = cases.make_pair(name)
    F0, v0 = _check(O, G, cfg, mask, strict)                 # nothing attached: the solar term only
    cases.attach(O, cfg, seaice=False)
    cases.attach(G, cfg, seaice=False)
    assert O.getdeps() == G.getdeps() and G.getdeps()["lvsc"] != 0.0
    F1, v1 = _check(O, G, cfg, mask, strict)                 # atmosphere
    assert not np.array_equal(F0, F1) and not np.array_equal(v0, v1)
    cases.attach(O, cfg, seaice=True)
    cases.attach(G, cfg, seaice=True)
    F2, v2 = _check(O, G, cfg, mask, strict)                 # atmosphere and sea ice
    assert not np.array_equal(F1, F2) and not np.array_equal(v1, v2)
    # a parameter change afterwards goes through the device forcing kernels (row N1 + forcing_cells_kernel)
    for par, val in (("COMB", 0.35), ("SUNP", 0.2), ("TEMP", 8.0)):
        O.set_par(par, val)
        G.setPar(par, val)
    _check(O, G, cfg, mask, strict)
    # new interface fields, as every Newton step of the coupled model brings them
    f = cases.coupling_fields(cfg, seed=11)
    for obj in (O, G):
        for k in ("tatm", "qatm", "albe"):
            obj.insert(k, f[k])
        obj.set_atmos_parameters(cases.ATMOS_PARS)
        obj.insert("msi", f["msi"])
        obj.set_seaice_parameters(cases.SEAICE_PARS)
    _check(O, G, cfg, mask, strict)


@pytest.mark.parametrize("name", ["coupled_small", "coupled_box"])
def test_coupling_blocks_and_their_products(name):
    import torch
    cfg, mask, O, G = cases.make_pair(name)
    f = cases.attach(O, cfg)
    cases.attach(G, cfg)
    n, m = cfg["n"], cfg["m"]
    x = cases.state(cfg, mask, seed=5)
    dx = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    dx[:G.nrows] = torch.from_numpy(x).cuda()
    assert np.array_equal(np.repeat(O.suno(), n), G.suno())
    # dF_T/dM_seaice: a kernel with the Fortran's expression order, no tanh -> the oracle's bits
    assert np.array_equal(O.dftdm(x), G.dftdm(dx))
    Ao, So = cases.blocks_from_oracle(O, cfg, mask, x, f, cases.ATMOS_PARS["da"])
    Ag, Sg = G.atmos_block(), G.seaice_block(dx)
    assert cases.compact(Ag) == Ao and cases.compact(Sg) == So
    assert len(Ao) > 0
    # CouplingBlock::applyMatrix on the device against a dense product in Epetra's order (entries of a row sorted by column)
    rng = np.random.default_rng(2)
    for blk, ref, ncols in ((Ag, Ao, 3 * n * m), (Sg, So, 4 * n * m)):
        C = G.coupling_block(*blk, ncols)
        v = rng.normal(size=ncols)
        dv = torch.from_numpy(v).cuda()
        out = torch.full((G.nvec,), 3.0, dtype=torch.float64, device="cuda")
        C.apply(dv, out)
        G.sync()
        want = np.zeros(G.nrows)
        for row, ent in ref:
            s = 0.0
            for c, a in sorted(ent):
                s = s + a * v[c]
            want[row] = s
        assert np.array_equal(out[:G.nrows].cpu().numpy(), want)
        # the ocean row of CoupledModel::applyMatrix (CoupledModel.C:436-470): out = J x + C v
        G.computeJacobian(dx)
        y = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        G.applyMatrix(dx, y)
        G.sync()
        base = y[:G.nrows].cpu().numpy().copy()
        C.apply(dv, y, accumulate=True)
        G.sync()
        assert np.array_equal(y[:G.nrows].cpu().numpy(), base + want)


def test_fortran_level_inserts_are_ignored_without_the_flag():
    """inserts.F90:24,43,62: an uncoupled ocean ignores atmosphere fields; the sea-ice mask is stored but lin does not read it"""
    cfg, mask, O, G = cases.make_pair("natl8")
    f = cases.coupling_fields(cfg)
    x = cases.state(cfg, mask, seed=5)
    F0 = G.computeRHS_host(x)
    for k in ("tatm", "qatm", "albe", "msi", "patm", "qsa", "gsi"):
        G.insert(k, f.get(k, f["tatm"]))
    G.set_atmos_parameters(cases.ATMOS_PARS)
    G.set_seaice_parameters(cases.SEAICE_PARS)
    assert np.array_equal(G.computeRHS_host(x), F0)
    assert np.all(G.atmos_block()[0].size == 0)
