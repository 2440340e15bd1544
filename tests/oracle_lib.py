"""ctypes binding of the CPU oracle (oracle/libthcm_oracle.so).  Test infrastructure only.

The oracle restates the reference's Fortran THCM kernels (see oracle/thcm_oracle.h); nothing in the
product package imports this module.
"""
import ctypes as C
import math
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_LIB = None

# par.F90:37-66 / THCM.C:1754-1869 (par2int)
PAR = dict(AL_T=1, RAYL=2, EK_V=3, EK_H=4, ROSB=5, MIXP=6, RESC=7, SPL1=8, HMTP=9, SUNP=10, PE_H=11,
           PE_V=12, P_VC=13, LAMB=14, SALT=15, WIND=16, TEMP=17, BIOT=18, COMB=19, ARCL=20, NLES=21,
           IFRICB=22, CONT=23, ENER=24, ALPC=25, CMPR=26, FPER=27, SPER=28, MKAP=29, SPL2=30)


class Config(C.Structure):
    _fields_ = [("n", C.c_int), ("m", C.c_int), ("l", C.c_int),
                ("xmin", C.c_double), ("xmax", C.c_double), ("ymin", C.c_double), ("ymax", C.c_double),
                ("periodic", C.c_int), ("hdim", C.c_double), ("qz", C.c_double),
                ("alphaT", C.c_double), ("alphaS", C.c_double),
                ("ih", C.c_int), ("vmix", C.c_int), ("tap", C.c_int), ("rho_mixing", C.c_int),
                ("coriolis_on", C.c_int), ("forcing_type", C.c_int),
                ("tres", C.c_int), ("sres", C.c_int), ("iza", C.c_int), ("ite", C.c_int), ("its", C.c_int),
                ("coupled_T", C.c_int), ("coupled_S", C.c_int),
                ("int_sign", C.c_int), ("intcond_i", C.c_int), ("intcond_j", C.c_int),
                ("fix_pressure_points", C.c_int)]


def default_config(n, m, l, xmin=286.0, xmax=350.0, ymin=10.0, ymax=74.0, periodic=False, hdim=4000.0,
                   qz=1.0, **kw):
    """THCM::getDefaultInitParameters (THCM.C:2748-2813); bounds in degrees like the XML files."""
    d = dict(n=n, m=m, l=l, xmin=xmin * math.pi / 180.0, xmax=xmax * math.pi / 180.0,
             ymin=ymin * math.pi / 180.0, ymax=ymax * math.pi / 180.0, periodic=int(periodic), hdim=hdim,
             qz=qz, alphaT=1.0e-4, alphaS=7.6e-4, ih=0, vmix=1, tap=1, rho_mixing=0, coriolis_on=1,
             forcing_type=0, tres=1, sres=1, iza=2, ite=1, its=1, coupled_T=0, coupled_S=0, int_sign=-1,
             intcond_i=-1, intcond_j=-1, fix_pressure_points=0)
    d.update(kw)
    return d


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ORACLE_DIR, "libthcm_oracle.so")
        src_newer = (not os.path.exists(path)) or any(
            os.path.getmtime(os.path.join(ORACLE_DIR, f)) > os.path.getmtime(path)
            for f in ("thcm_oracle.cpp", "gmres_oracle.cpp", "thcm_oracle.h"))
        if src_newer:
            build()
        L = C.CDLL(path)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.thcm_oracle_create.restype = C.c_void_p
        L.thcm_oracle_create.argtypes = [C.POINTER(Config), ip]
        L.thcm_oracle_destroy.argtypes = [C.c_void_p]
        L.thcm_oracle_set_par.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.thcm_oracle_get_par.argtypes = [C.c_void_p, C.c_int]
        L.thcm_oracle_get_par.restype = C.c_double
        L.thcm_oracle_ndim.argtypes = [C.c_void_p]
        L.thcm_oracle_fortran_rhs.argtypes = [C.c_void_p, dp, dp]
        L.thcm_oracle_fortran_matrix.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_fortran_nnz.argtypes = [C.c_void_p]
        L.thcm_oracle_fortran_nnz.restype = C.c_int64
        L.thcm_oracle_fortran_csr.argtypes = [C.c_void_p, ip, ip, dp, dp]
        L.thcm_oracle_frc.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_graph_nnz.argtypes = [C.c_void_p]
        L.thcm_oracle_graph_nnz.restype = C.c_int64
        L.thcm_oracle_graph.argtypes = [C.c_void_p, ip, ip]
        L.thcm_oracle_eval_rhs.argtypes = [C.c_void_p, dp, dp]
        L.thcm_oracle_eval_jac.argtypes = [C.c_void_p, dp, dp, dp]
        L.thcm_oracle_rowintcon.argtypes = [C.c_void_p]
        L.thcm_oracle_intcond_coeff.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_set_intcond_correction.argtypes = [C.c_void_p, C.c_double]
        L.thcm_oracle_dropped_entries.argtypes = [C.c_void_p]
        L.thcm_oracle_dropped_entries.restype = C.c_int64
        L.thcm_oracle_set_atmos_parameters.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_set_seaice_parameters.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_insert.argtypes = [C.c_void_p, C.c_int, dp]
        L.thcm_oracle_getdeps.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_suno.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_dftdm.argtypes = [C.c_void_p, dp, dp]
        L.thcm_oracle_average_block.argtypes = [C.c_void_p, dp]
        L.thcm_oracle_scal.argtypes = [dp, dp, dp]
        L.thcm_oracle_recompute_scaling.argtypes = [C.c_void_p, dp, dp]
        L.thcm_oracle_spmv.argtypes = [C.c_int, ip, ip, dp, dp, dp]
        L.thcm_oracle_gmres.argtypes = [C.c_int, ip, ip, dp, dp, dp, C.c_int, C.c_int, C.c_double, C.c_int, dp, dp, ip]
        L.thcm_oracle_gmres.restype = C.c_int
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


class Oracle:
    def __init__(self, cfg, landm):
        self.cfg = dict(cfg)
        self.c = Config(**cfg)
        self.landm = np.ascontiguousarray(landm, dtype=np.int32).ravel()
        assert self.landm.size == (cfg["n"] + 2) * (cfg["m"] + 2) * (cfg["l"] + 2)
        self.h = lib().thcm_oracle_create(C.byref(self.c), _ip(self.landm))
        if not self.h:
            raise RuntimeError("thcm_oracle_create failed")
        self.ndim = lib().thcm_oracle_ndim(self.h)
        self._graph = None

    def __del__(self):
        if getattr(self, "h", None):
            lib().thcm_oracle_destroy(self.h)
            self.h = None

    def set_par(self, name, value):
        lib().thcm_oracle_set_par(self.h, PAR[name] if isinstance(name, str) else name, float(value))

    def get_par(self, name):
        return lib().thcm_oracle_get_par(self.h, PAR[name] if isinstance(name, str) else name)

    # ---- coupling: usrc.F90:237-333, inserts.F90, getdeps, probe.F90:381-444 ----
    FIELDS = dict(tatm=0, qatm=1, albe=2, patm=3, qsa=4, msi=5, gsi=6)
    ATMOS_PARS = ("tdim", "qdim", "nuq", "eta", "dqso", "dqsi", "dqdt", "Eo0", "Ei0", "Cs", "t0o", "t0i", "a0", "da", "tauf", "tauc", "comb", "albf")
    SEAICE_PARS = ("zeta", "a0", "Lf", "s0", "rhoo", "qvar", "q0")

    def set_atmos_parameters(self, pars):
        a = np.array([float(pars[k]) for k in self.ATMOS_PARS])
        lib().thcm_oracle_set_atmos_parameters(self.h, _dp(a))

    def set_seaice_parameters(self, pars):
        a = np.array([float(pars[k]) for k in self.SEAICE_PARS])
        lib().thcm_oracle_set_seaice_parameters(self.h, _dp(a))

    def insert(self, which, field):
        f = np.ascontiguousarray(field, np.float64).ravel()
        assert f.size == self.cfg["n"] * self.cfg["m"]
        lib().thcm_oracle_insert(self.h, self.FIELDS[which], _dp(f))

    def getdeps(self):
        out = np.zeros(7)
        lib().thcm_oracle_getdeps(self.h, _dp(out))
        return dict(zip(("Ooa", "Os", "nus", "eta", "lvsc", "qdim", "pQSnd"), out))

    def suno(self):
        out = np.zeros(self.cfg["m"])
        lib().thcm_oracle_suno(self.h, _dp(out))
        return out

    def dftdm(self, x):
        x = np.ascontiguousarray(x, np.float64)
        out = np.zeros(self.cfg["n"] * self.cfg["m"])
        lib().thcm_oracle_dftdm(self.h, _dp(x), _dp(out))
        return out

    def graph(self):
        if self._graph is None:
            nnz = lib().thcm_oracle_graph_nnz(self.h)
            rp = np.zeros(self.ndim + 1, np.int32)
            col = np.zeros(nnz, np.int32)
            lib().thcm_oracle_graph(self.h, _ip(rp), _ip(col))
            self._graph = (rp, col)
        return self._graph

    def rhs(self, x):
        x = np.ascontiguousarray(x, np.float64)
        F = np.zeros(self.ndim)
        lib().thcm_oracle_eval_rhs(self.h, _dp(x), _dp(F))
        return F

    def jac(self, x):
        x = np.ascontiguousarray(x, np.float64)
        rp, col = self.graph()
        val = np.zeros(col.size)
        diagB = np.zeros(self.ndim)
        lib().thcm_oracle_eval_jac(self.h, _dp(x), _dp(val), _dp(diagB))
        return val, diagB

    def dropped(self):
        return lib().thcm_oracle_dropped_entries(self.h)

    def fortran_rhs(self, x):
        x = np.ascontiguousarray(x, np.float64)
        B = np.zeros(self.ndim)
        lib().thcm_oracle_fortran_rhs(self.h, _dp(x), _dp(B))
        return B

    def fortran_matrix(self, x):
        x = np.ascontiguousarray(x, np.float64)
        lib().thcm_oracle_fortran_matrix(self.h, _dp(x))
        nnz = lib().thcm_oracle_fortran_nnz(self.h)
        beg = np.zeros(self.ndim + 1, np.int32)
        jco = np.zeros(nnz, np.int32)
        co = np.zeros(nnz)
        coB = np.zeros(self.ndim)
        lib().thcm_oracle_fortran_csr(self.h, _ip(beg), _ip(jco), _dp(co), _dp(coB))
        return beg, jco, co, coB

    def frc(self):
        f = np.zeros(self.ndim)
        lib().thcm_oracle_frc(self.h, _dp(f))
        return f

    def average_block(self):
        """m_scaling::average_block on the Fortran CSR of the last jac()/fortran_matrix() call: db[ii, jj]"""
        db = np.zeros(36)
        lib().thcm_oracle_average_block(self.h, _dp(db))
        return db.reshape(6, 6).T.copy()

    def recompute_scaling(self):
        """THCM::RecomputeScaling: (rowScaling, colScaling) or None if the averaged block is singular"""
        r, c = np.zeros(self.ndim), np.zeros(self.ndim)
        return (r, c) if lib().thcm_oracle_recompute_scaling(self.h, _dp(r), _dp(c)) else None

    def rowintcon(self):
        return lib().thcm_oracle_rowintcon(self.h)

    def intcond_coeff(self):
        c = np.zeros(self.ndim)
        lib().thcm_oracle_intcond_coeff(self.h, _dp(c))
        return c


def scal(db):
    """m_scaling::scal on a 6x6 array db[ii, jj]: (dr, dc) or None if singular"""
    a = np.ascontiguousarray(np.asarray(db, np.float64).T).ravel()  # column-major like the Fortran array
    dr, dc = np.zeros(6), np.zeros(6)
    return (dr, dc) if lib().thcm_oracle_scal(_dp(a), _dp(dr), _dp(dc)) else None


def spmv(rowptr, col, val, x):
    y = np.zeros(rowptr.size - 1)
    x = np.ascontiguousarray(x, np.float64)
    lib().thcm_oracle_spmv(rowptr.size - 1, _ip(rowptr), _ip(col), _dp(val), _dp(x), _dp(y))
    return y


def gmres(rowptr, col, val, b, x0, m, maxit, tol, blockinv=None):
    n = rowptr.size - 1
    x = np.array(x0, np.float64)
    hist = np.zeros(maxit + m + 4)
    iters = C.c_int(0)
    bi = _dp(np.ascontiguousarray(blockinv)) if blockinv is not None else None
    rc = lib().thcm_oracle_gmres(n, _ip(rowptr), _ip(col), _dp(val), _dp(np.ascontiguousarray(b, np.float64)), _dp(x),
                                 m, maxit, tol, 0 if blockinv is None else 1, bi, _dp(hist), C.byref(iters))
    return rc, x, hist, iters.value


# ---------------------------------------------------------------------------- masks
def read_mask(path, n, m, l):
    """topo.F90:41-75 (readmask): per level one header line then rows j=m+1..0, each n+2 digits."""
    lines = open(path).read().split("\n")
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    p = 0
    for k in range(l + 2):
        p += 1
        for j in range(m + 1, -1, -1):
            a[k, j, :] = [int(ch) for ch in lines[p][:n + 2]]
            p += 1
    # land inversion fix, topo.F90:94-103
    for k in range(l, 1, -1):
        inv = (a[k, 1:m + 1, 1:n + 1] == 1) & (a[k - 1, 1:m + 1, 1:n + 1] == 0)
        a[k - 1, 1:m + 1, 1:n + 1][inv] = 1
    return a


def no_land_mask(n, m, l, periodic):
    """depth3land itopo=1 (topo.F90:167-177, :293-298)."""
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    a[1:l + 1, 1:m + 1, 1:n + 1] = 0
    if periodic:
        sel = (a[:, :, 1] == 0) & (a[:, :, n] == 0)
        a[:, :, 0][sel] = 3
        a[:, :, n + 1][sel] = 3
    return a


def random_mask(n, m, l, periodic, seed, land_frac=0.25):
    """Synthetic bathymetry: random column depths (no land above ocean), reference border convention."""
    rng = np.random.default_rng(seed)
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    depth = rng.integers(0, l + 1, size=(m, n))          # number of ocean layers from the top
    depth[rng.random((m, n)) > land_frac * 2] = l        # most columns full depth
    depth[rng.random((m, n)) < land_frac / 2] = 0        # some columns all land
    for k in range(1, l + 1):
        a[k, 1:m + 1, 1:n + 1] = np.where(k > l - depth, 0, 1)
    if periodic:
        sel = (a[:, :, 1] == 0) & (a[:, :, n] == 0)
        a[:, :, 0] = np.where(sel, 3, 1)
        a[:, :, n + 1] = np.where(sel, 3, 1)
    return a
