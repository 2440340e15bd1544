"""Shared parity cases: (name, config dict, mask, starting parameters)."""
import numpy as np

from iemic_b200 import grids


def _apply(obj, pars):
    for k, v in pars.items():
        obj.set_par(k, v) if hasattr(obj, "set_par") else obj.setPar(k, v)


def case(name):
    g = grids
    if name == "natl8":        # test/ocean: 8x8x4 North Atlantic box, mask_natl8, SRES=0 (integral condition)
        cfg = g.default_config(8, 8, 4, sres=0)
        return cfg, g.fixture_mask("mask_natl8"), dict(COMB=0.1, SALT=1.0, WIND=1.0, TEMP=10.0)
    if name == "natl8_sres1":
        cfg = g.default_config(8, 8, 4, sres=1)
        return cfg, g.fixture_mask("mask_natl8"), dict(COMB=0.1, SALT=1.0, WIND=1.0, TEMP=10.0)
    if name == "test6x6x4":    # test/matrix mask
        cfg = g.default_config(6, 6, 4, xmin=286, xmax=350, ymin=10, ymax=74, periodic=False)
        return cfg, g.fixture_mask("test6x6x4"), dict(COMB=0.1, SALT=1.0, WIND=1.0, TEMP=10.0)
    if name == "2dmoc":        # test/2dmoc: 3x6x6 periodic
        cfg = g.moc2d_config(3, 6, 6)
        return cfg, g.no_land_mask(3, 6, 6, True), dict(g.STARTING_PARS["2dmoc"], COMB=0.5, SALT=0.2)
    if name == "2dmoc_run":    # run/2dmoc: 4x32x16
        cfg = g.moc2d_config(4, 32, 16)
        return cfg, g.no_land_mask(4, 32, 16, True), dict(g.STARTING_PARS["2dmoc"], COMB=0.5, SALT=0.2)
    if name == "2dmoc_sres1":  # periodic in x inside ONE rank with the ranks stacked in y (seam rows next to halo rows)
        cfg = g.moc2d_config(3, 6, 6, sres=1)
        return cfg, g.no_land_mask(3, 6, 6, True), dict(g.STARTING_PARS["2dmoc"], COMB=0.5, SALT=0.2)
    if name == "2dmoc_run_sres1":
        cfg = g.moc2d_config(4, 32, 16, sres=1)
        return cfg, g.no_land_mask(4, 32, 16, True), dict(g.STARTING_PARS["2dmoc"], COMB=0.5, SALT=0.2)
    if name == "rand_periodic":
        cfg = g.global_config(37, 11, 7, ymin=-60, ymax=60)
        return cfg, g.random_mask(37, 11, 7, True, 5), dict(g.STARTING_PARS["global"])
    if name == "rand_box":
        cfg = g.default_config(35, 9, 6, ymin=-40, ymax=50, qz=1.8)
        return cfg, g.random_mask(35, 9, 6, False, 7), dict(g.STARTING_PARS["global"])
    if name == "rand_fixp":    # pressure points fixed + rho mixing
        cfg = g.default_config(12, 7, 5, fix_pressure_points=1, rho_mixing=1)
        return cfg, g.random_mask(12, 7, 5, False, 9, land_frac=0.1), dict(g.STARTING_PARS["global"])
    if name == "global4":      # BASELINE config 2: global 4 degree, 96x38x12
        cfg = g.global_config(96, 38, 12)
        return cfg, g.synthetic_mask(96, 38, 12), dict(g.STARTING_PARS["global"])
    if name == "coupled4":     # BASELINE config 3 (ocean block): Coupled Temperature=1, qz=1.5
        cfg = g.global_config(96, 38, 12, qz=1.5, coupled_T=1)
        return cfg, g.synthetic_mask(96, 38, 12), dict(g.STARTING_PARS["global"])
    if name == "coupled_small":  # row N2: small periodic globe with land, Coupled Temperature = 1, non-zero solar forcing
        cfg = g.global_config(24, 12, 6, qz=1.5, coupled_T=1)
        return cfg, g.random_mask(24, 12, 6, True, 13), dict(g.STARTING_PARS["global"], SUNP=0.7)
    if name == "coupled_small_nomix":  # Mixing = 0: no tanh anywhere, the CUDA path must reproduce every bit of the oracle
        cfg = g.global_config(24, 12, 6, qz=1.5, coupled_T=1, vmix=0)
        return cfg, g.random_mask(24, 12, 6, True, 13), dict(g.STARTING_PARS["global"], SUNP=0.7)
    if name == "coupled_box":    # same without periodicity, pressure points fixed
        cfg = g.default_config(14, 9, 5, coupled_T=1, fix_pressure_points=1)
        return cfg, g.random_mask(14, 9, 5, False, 17, land_frac=0.15), dict(g.STARTING_PARS["global"], SUNP=1.0)
    if name == "global_half":  # 48x19x12 sweep point
        cfg = g.global_config(48, 19, 12)
        return cfg, g.upsample_mask(g.fixture_mask("mask_global_96x38x12"), 48, 19, 12), dict(g.STARTING_PARS["global"])
    raise KeyError(name)


def reference_state():
    """The reference's own golden state (tests/golden/ocean_reference.npz, from test/ocean/ocean_reference.h5): the end of
    the continuation of src/tests/reft_ocean.C on the 16^3 periodic gateway box (test/ocean/reft_ocean_params.xml).
    Returns (config, mask, {parameter index: value}, state).  "Mixing" = 2 there: convective adjustment switched on as
    soon as T and S are non-zero (vmix_control, mix_imp.f:131-166), which is Mixing = 1 at this state."""
    import os
    from iemic_b200.capi import PAR_NAMES
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ocean_reference.npz"))
    cfg = grids.default_config(16, 16, 16, xmin=300.0, xmax=340.0, ymin=20.0, ymax=60.0, periodic=True, hdim=4000.0, qz=1.0,
                               forcing_type=2, tres=1, sres=1, vmix=1)
    idx = dict(PAR_NAMES, FPER=27)
    pars = {idx[str(n)]: float(v) for n, v in zip(z["par_names"], z["par_values"])}
    return cfg, z["mask"].astype(np.int32), pars, z["state"].astype(np.float64)


def state(cfg, mask, seed=1, kind="random"):
    return grids.synthetic_state(cfg["n"], cfg["m"], cfg["l"], mask, seed=seed, kind=kind)


def make_pair(name):
    """(oracle, product Ocean) with identical parameters"""
    import oracle_lib as ol
    import iemic_b200 as ib
    cfg, mask, pars = case(name)
    O = ol.Oracle(cfg, mask)
    G = ib.Ocean(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        G.setPar(k, v)
    return cfg, mask, O, G


# ---- row N2: an attached atmosphere and sea ice (values of the order the reference's models hand over) ----
ATMOS_PARS = dict(tdim=1.0, qdim=0.01, nuq=2.58e-2, eta=4.68e-2, dqso=5.6e-4, dqsi=6.4e-4, dqdt=5.0e-4, Eo0=3.1e-8, Ei0=2.2e-8, Cs=1.0e-3,
                  t0o=15.0, t0i=-5.0, a0=0.3, da=0.5, tauf=1.4e5, tauc=3.0e6, comb=1.0, albf=1.0)
SEAICE_PARS = dict(zeta=1.7e3, a0=-0.0575, Lf=3.347e5, s0=35.0, rhoo=1.024e3, qvar=5.0e-1, q0=-1.0e2)


def coupling_fields(cfg, seed=3):
    """synthetic interface fields on the n x m surface (i fastest): atmosphere temperature / humidity / albedo anomalies and a
    sea-ice mask with open water, full cover and fractional values (the reference's mask is a smooth tanh of the ice thickness)"""
    n, m = cfg["n"], cfg["m"]
    rng = np.random.default_rng(seed)
    lat = np.repeat(np.linspace(-1.0, 1.0, m), n)
    f = dict(tatm=rng.normal(0.0, 3.0, n * m) - 10.0 * lat ** 2,
             qatm=rng.normal(0.0, 2.0e-3, n * m),
             albe=rng.uniform(0.0, 1.0, n * m),
             msi=np.where(np.abs(lat) > 0.7, 1.0, np.where(np.abs(lat) > 0.45, rng.uniform(0.0, 1.0, n * m), 0.0)))
    return f


def attach(obj, cfg, seed=3, seaice=True):
    """Ocean::synchronize(atmos) then Ocean::synchronize(seaice) (Ocean.C:1443-1493): inserts, then the parameter calls"""
    f = coupling_fields(cfg, seed)
    for k in ("tatm", "qatm", "albe"):
        obj.insert(k, f[k])
    obj.set_atmos_parameters(ATMOS_PARS)
    if seaice:
        obj.insert("msi", f["msi"])
        obj.set_seaice_parameters(SEAICE_PARS)
    return f


def blocks_from_oracle(O, cfg, mask, x, f, albed):
    """Ocean::getBlock(Atmosphere) / getBlock(SeaIce), T rows, restated on the oracle's numbers (Ocean.C:1538-1746)"""
    n, m, l = cfg["n"], cfg["m"], cfg["l"]
    d = O.getdeps()
    comb, sunp = O.get_par("COMB"), O.get_par("SUNP")
    suno, dftdm = O.suno(), O.dftdm(x)
    surf = np.asarray(mask).reshape(l + 2, m + 2, n + 2)[l, 1:m + 1, 1:n + 1]
    A, S = [], []
    for j in range(m):
        for i in range(n):
            if surf[j, i] != 0:
                continue
            sr = j * n + i
            M = f["msi"][sr]
            row = 6 * ((l - 1) * n * m + n * j + i) + 4
            dTFT = d["Ooa"] * (1.0 - M)
            dAFT = -comb * sunp * suno[j] * albed * (1.0 - M)
            dQFT = d["lvsc"] * d["eta"] * d["qdim"] * (1.0 - M)
            A.append((row, [(3 * sr + 0, -dTFT), (3 * sr + 2, -dAFT), (3 * sr + 1, -dQFT)]))
            S.append((row, [(4 * sr + 2, -dftdm[sr])]))
    return A, S


def compact(block):
    rows, beg, jco, co = block
    return [(int(rows[r]), [(int(jco[p]), float(co[p])) for p in range(beg[r], beg[r + 1])]) for r in range(rows.size)]
