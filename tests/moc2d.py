"""The reference's 2-D MOC known-answer test (src/tests/intt_2dmoc.C:15-137) as a driver over any backend that offers
rhs(x), jac(x) -> (values, diagB), set_par, get_par, graph().  Natural-parameter continuation with Newton and a dense
least-squares solve (648 unknowns) stands in for Continuation + Belos/MRILU: the fixed points do not depend on the solver."""
import numpy as np
import scipy.sparse as sp

from iemic_b200 import grids

N_, M_, L_ = 3, 6, 6


def setup(make_backend):
    cfg = grids.moc2d_config(N_, M_, L_)
    mask = grids.no_land_mask(N_, M_, L_, True)
    B = make_backend(cfg, mask)
    for k, v in grids.STARTING_PARS["2dmoc"].items():
        B.set_par(k, v)
    return cfg, mask, B


def newton(B, x, rp, col, tol=1e-9, maxit=15):
    n = x.size
    for _ in range(maxit):
        F = B.rhs(x)
        if np.linalg.norm(F) < tol:
            return x, True
        val, _ = B.jac(x)
        J = sp.csr_matrix((val, col, rp), shape=(n, n)).toarray()
        x = x + np.linalg.lstsq(J, -F, rcond=1e-12)[0]   # pressure modes make J rank deficient: minimum-norm update
    return x, np.linalg.norm(B.rhs(x)) < tol


def continuation(B, x, rp, col, par, target, step):
    v = B.get_par(par)
    while abs(v - target) > 1e-14:
        vn = v + step
        if (step > 0 and vn > target) or (step < 0 and vn < target):
            vn = target
        B.set_par(par, vn)
        xn, ok = newton(B, x.copy(), rp, col)
        if not ok:
            B.set_par(par, v)
            step /= 2
            assert abs(step) > 1e-6, "continuation stuck at %s = %g" % (par, v)
            continue
        x, v = xn, vn
    return x


def psim(cfg, x):
    """Ocean::getPsiM (src/ocean/Ocean.C:872-888) = OceanGrid::recomputePsiM (OceanGrid.C:270-345) + compute_psim
    (src/ocean/thcm_utils.F90:95-118): meridional overturning streamfunction in Sv, (min, max)."""
    n, m, l = cfg["n"], cfg["m"], cfg["l"]
    hdim, r0dim, udim = cfg["hdim"], 6.37e6, 0.1
    dx = (cfg["xmax"] - cfg["xmin"]) / n
    dy = (cfg["ymax"] - cfg["ymin"]) / m
    dz = 1.0 / l
    qz = cfg["qz"]
    assert qz <= 1.0
    ze = (np.arange(1, l + 1) - 0.5) * dz - 1.0
    z = ze + (1 - qz) * ze * (1 - ze)
    dfzT = 1.0 + (1 - qz) * (1 - 2 * ze)
    yv = np.arange(0, m + 1) * dy + cfg["ymin"]
    X = x.reshape(l, m, n, 6)
    psi = np.zeros((m + 1, l + 1))
    for j in range(1, m + 1):
        for k in range(1, l + 1):
            vs = X[k - 1, j - 1, :, 1].sum() * dx
            psi[j, k] = psi[j, k - 1]
            if z[k - 1] * hdim < -500:
                psi[j, k] = -np.cos(yv[j]) * vs * dz * dfzT[k - 1] + psi[j, k - 1]
    sv = r0dim * hdim * udim * 1e-6
    return psi.min() * sv, psi.max() * sv


def run_branch(make_backend, sign):
    """Stages of intt_2dmoc.C: forcing 0->1, asymmetry CMPR 0 -> -/+0.2, salinity forcing 0 -> 0.02, CMPR -> 0, salinity
    forcing -> 0.04.  sign=-1 is the reference's path to 'state 1'; sign=+1 reaches the mirror image ('state 2')."""
    cfg, mask, B = setup(make_backend)
    rp, col = B.graph()
    x = np.zeros(rp.size - 1)
    x = continuation(B, x, rp, col, "COMB", 1.0, 0.1)
    sym = psim(cfg, x)
    x = continuation(B, x, rp, col, "CMPR", sign * 0.2, sign * 0.05)
    x = continuation(B, x, rp, col, "SALT", 0.02, 0.005)
    x = continuation(B, x, rp, col, "CMPR", 0.0, -sign * 0.05)
    x = continuation(B, x, rp, col, "SALT", 0.04, 0.005)
    return sym, psim(cfg, x), x
