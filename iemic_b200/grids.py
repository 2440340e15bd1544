"""Synthetic workloads of the hot path: grid configurations, land masks and states.

Only numpy; nothing here touches the reference tree at run time.  The 4-degree global land mask and the small
test masks are committed fixtures (tests/golden/masks.npz, made by tests/golden/make_fixtures.py).
"""
import math
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MASKS_NPZ = os.path.join(ROOT, "tests", "golden", "masks.npz")


def default_config(n, m, l, xmin=286.0, xmax=350.0, ymin=10.0, ymax=74.0, periodic=False, hdim=4000.0,
                   qz=1.0, **kw):
    """THCM::getDefaultInitParameters (src/ocean/THCM.C:2748-2813); bounds in degrees like the XML files."""
    d = dict(n=n, m=m, l=l, xmin=xmin * math.pi / 180.0, xmax=xmax * math.pi / 180.0,
             ymin=ymin * math.pi / 180.0, ymax=ymax * math.pi / 180.0, periodic=int(periodic), hdim=hdim,
             qz=qz, alphaT=1.0e-4, alphaS=7.6e-4, ih=0, vmix=1, tap=1, rho_mixing=0, coriolis_on=1,
             forcing_type=0, tres=1, sres=1, iza=2, ite=1, its=1, coupled_T=0, coupled_S=0, int_sign=-1,
             intcond_i=-1, intcond_j=-1, fix_pressure_points=0)
    d.update(kw)
    return d


def global_config(n, m, l, qz=2.25, **kw):
    """run/ocean/global/ocean_params.xml: periodic globe, 85.5S-85.5N, 5000 m, stretched grid."""
    d = dict(xmin=0.0, xmax=360.0, ymin=-85.5, ymax=85.5, periodic=True, hdim=5000.0, qz=qz)
    d.update(kw)
    return default_config(n, m, l, **d)


def moc2d_config(n=3, m=6, l=6, **kw):
    """test/2dmoc/ocean_params.xml (run/2dmoc uses 4x32x16): periodic lat-depth box, no Coriolis, forcing type 1,
    salinity flux forcing with the integral condition."""
    d = dict(xmin=286.0, xmax=350.0, ymin=-60.0, ymax=60.0, periodic=True, hdim=4000.0, qz=1.0, coriolis_on=0, forcing_type=1, sres=0)
    d.update(kw)
    return default_config(n, m, l, **d)


# parameters set by the "Starting Parameters" lists
STARTING_PARS = {
    "global": dict(COMB=0.1, SALT=1.0, WIND=1.0, TEMP=10.0, SPL1=2.0e3, SPL2=0.01),
    "2dmoc": dict(COMB=0.0, SUNP=0.0, SALT=0.0, WIND=0.0, TEMP=10.0, SPL1=2.0e3, SPL2=0.01, P_VC=0.0, ROSB=0.0, CMPR=0.0,
                  EK_H=371.764, RAYL=15.6869),
}


def fix_mask(a):
    """land inversion fix of topo.F90:94-103 on an (l+2, m+2, n+2) mask"""
    l, m, n = a.shape[0] - 2, a.shape[1] - 2, a.shape[2] - 2
    for k in range(l, 1, -1):
        inv = (a[k, 1:m + 1, 1:n + 1] == 1) & (a[k - 1, 1:m + 1, 1:n + 1] == 0)
        a[k - 1, 1:m + 1, 1:n + 1][inv] = 1
    return a


def no_land_mask(n, m, l, periodic):
    """depth3land itopo=1 (topo.F90:167-177, :293-298)."""
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    a[1:l + 1, 1:m + 1, 1:n + 1] = 0
    _periodic_ring(a, periodic)
    return a


def _periodic_ring(a, periodic):
    n = a.shape[2] - 2
    if periodic:
        sel = (a[:, :, 1] == 0) & (a[:, :, n] == 0)
        a[:, :, 0] = np.where(sel, 3, 1)
        a[:, :, n + 1] = np.where(sel, 3, 1)
    else:
        a[:, :, 0] = 1
        a[:, :, n + 1] = 1


def random_mask(n, m, l, periodic, seed, land_frac=0.25):
    """Synthetic bathymetry: random column depths (no land above ocean), reference border convention."""
    rng = np.random.default_rng(seed)
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    depth = rng.integers(0, l + 1, size=(m, n))          # number of ocean layers from the top
    depth[rng.random((m, n)) > land_frac * 2] = l        # most columns full depth
    depth[rng.random((m, n)) < land_frac / 2] = 0        # some columns all land
    for k in range(1, l + 1):
        a[k, 1:m + 1, 1:n + 1] = np.where(k > l - depth, 0, 1)
    _periodic_ring(a, periodic)
    return a


def fixture_mask(name):
    """A committed mask fixture (tests/golden/masks.npz): e.g. 'mask_global_96x38x12', 'mask_natl8', 'test6x6x4'."""
    with np.load(MASKS_NPZ) as z:
        return z[name].astype(np.int32)


def upsample_mask(src, n, m, l, periodic=True):
    """Nearest-neighbour up-sampling of an (ls+2, ms+2, ns+2) mask to n x m x l interior cells, then the land inversion
    fix.  This is the synthetic 1-degree mask of BASELINE config 4 (4-degree mask x (3.75, 4.2, 2.67))."""
    ls, ms, ns = src.shape[0] - 2, src.shape[1] - 2, src.shape[2] - 2
    ki = np.minimum((np.arange(l) * ls) // l, ls - 1) + 1
    ji = np.minimum((np.arange(m) * ms) // m, ms - 1) + 1
    ii = np.minimum((np.arange(n) * ns) // n, ns - 1) + 1
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    inner = src[np.ix_(ki, ji, ii)]
    a[1:l + 1, 1:m + 1, 1:n + 1] = np.where(inner == 0, 0, 1)
    fix_mask(a)
    _periodic_ring(a, periodic)
    return a


def synthetic_mask(n, m, l, periodic=True, kind="global"):
    if kind == "none":
        return no_land_mask(n, m, l, periodic)
    src = fixture_mask("mask_global_96x38x12")
    if (n, m, l) == (96, 38, 12):
        a = src.copy()
        fix_mask(a)
        return a
    return upsample_mask(src, n, m, l, periodic)


def synthetic_state(n, m, l, mask, seed=20261104, kind="random"):
    """x = 0.1 * s_var * U(-1,1) generated by GLOBAL index (decomposition independent), land cells zeroed
    (Ocean::applyLandMask semantics), surface w = 0.  kind='smooth': T = cos(pi y / ymax) (z+1), S = T/2."""
    ncell = n * m * l
    if kind == "random":
        rng = np.random.default_rng(seed)
        x = 0.1 * rng.uniform(-1.0, 1.0, size=(ncell, 6))
        x[:, 2] *= 1.0e-2
    else:
        x = np.zeros((ncell, 6))
        jj = (np.arange(m) + 0.5) / m * 2.0 - 1.0
        kk = (np.arange(l) + 0.5) / l
        T = np.cos(np.pi * jj)[None, :, None] * kk[:, None, None] * np.ones((1, 1, n))
        x[:, 4] = T.ravel()
        x[:, 5] = 0.5 * T.ravel()
    land = (mask[1:l + 1, 1:m + 1, 1:n + 1] != 0).ravel()
    x[land, :] = 0.0
    x3 = x.reshape(l, m, n, 6)
    x3[l - 1, :, :, 2] = 0.0
    return x.ravel()


def decomp2d(n, m, nprocs):
    """TRIOS::Domain::Decomp2D (src/trios/TRIOS_Domain.C:81-135): returns (npN, npM)."""
    npN, npM = 1, nprocs
    r_min = 100
    for t1 in range(nprocs, 0, -1):
        t2 = nprocs // t1
        r = abs(m // t1 - n // t2)
        if t1 * t2 == nprocs and r <= r_min:
            r_min = r
            npM, npN = t1, t2
    return npN, npM
