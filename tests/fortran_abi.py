"""ctypes driver for seam S1 (libthcm_b200_fortran.so): calls the Fortran-named entry points the way THCM.C does
(src/ocean/THCM.C:189-800 constructor sequence, :949-1192 evaluate)."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "iemic_b200", "lib", "libthcm_b200_fortran.so")
HEADER = os.path.join(ROOT, "include", "thcm_fortran_abi.h")


def declared():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\bvoid\s+([A-Za-z_][A-Za-z0-9_]*)\s*\(", src)))


def load():
    return C.CDLL(LIB)


def _i(v):
    return C.byref(C.c_int(int(v)))


def _d(v):
    return C.byref(C.c_double(float(v)))


class FortranTHCM:
    """One process-global model, like the Fortran module state."""

    def __init__(self, cfg, landm):
        self.L = load()
        self.cfg = cfg
        n, m, l = cfg["n"], cfg["m"], cfg["l"]
        e = b"\0"
        getattr(self.L, "__m_global_MOD_initialize")(_i(n), _i(m), _i(l), _d(cfg["xmin"]), _d(cfg["xmax"]), _d(cfg["ymin"]), _d(cfg["ymax"]),
                                         _d(cfg["hdim"]), _d(cfg["qz"]), _i(cfg["periodic"]), _i(0), _i(0), _i(0),
                                         _i(cfg["tres"]), _i(cfg["sres"]), _i(cfg["iza"]), _i(cfg["ite"]), _i(cfg["its"]), _i(0),
                                         _i(cfg["coupled_T"]), _i(cfg["coupled_S"]), _i(cfg["forcing_type"]), e, e, e, e, e)
        lm = np.ascontiguousarray(landm, np.int32).ravel()
        z = np.zeros(n * m)
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        self.L.init_(_i(n), _i(m), _i(l), _i(n * m * l), _d(cfg["xmin"]), _d(cfg["xmax"]), _d(cfg["ymin"]), _d(cfg["ymax"]),
                     _d(cfg["alphaT"]), _d(cfg["alphaS"]), _i(cfg["ih"]), _i(cfg["vmix"]), _i(cfg["tap"]), _i(cfg["rho_mixing"]),
                     _i(cfg["coriolis_on"]), _i(cfg["periodic"]), lm.ctypes.data_as(C.POINTER(C.c_int)), dp(z), dp(z), dp(z), dp(z), dp(z))
        nrows, nnz = C.c_int(0), C.c_int(0)
        getattr(self.L, "__m_mat_MOD_get_array_sizes")(C.byref(nrows), C.byref(nnz))
        self.nrows, self.cap = nrows.value, nnz.value
        # THCM.C:631-650: the caller owns the CSR arrays
        self.begA = np.zeros(self.nrows + 1, np.int32); self.jcoA = np.zeros(self.cap, np.int32); self.coA = np.zeros(self.cap)
        self.coB = np.zeros(self.nrows)
        self.begF = np.zeros(self.nrows + 1, np.int32); self.jcoF = np.zeros(1, np.int32); self.coF = np.zeros(1)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
        getattr(self.L, "__m_mat_MOD_set_pointers")(C.byref(nrows), C.byref(nnz), ip(self.begA), ip(self.jcoA), dp(self.coA), dp(self.coB),
                                        ip(self.begF), ip(self.jcoF), dp(self.coF))

    def setpar(self, idx1, value):
        self.L.setparcs_(_i(idx1), _d(value))

    def getpar(self, idx1):
        v = C.c_double(0.0)
        self.L.getparcs_(_i(idx1), C.byref(v))
        return v.value

    def rhs(self, un):
        un = np.ascontiguousarray(un, np.float64)
        b = np.zeros(self.nrows)
        self.L.rhs_(un.ctypes.data_as(C.POINTER(C.c_double)), b.ctypes.data_as(C.POINTER(C.c_double)))
        return b

    def matrix(self, un):
        un = np.ascontiguousarray(un, np.float64)
        self.L.matrix_(un.ctypes.data_as(C.POINTER(C.c_double)))
        nnz = self.begA[self.nrows] - 1
        return self.begA.copy(), self.jcoA[:nnz].copy(), self.coA[:nnz].copy(), self.coB.copy()

    def setsres(self, sres):
        self.L.setsres_(_i(sres))

    def intcond_scaling(self):
        ncell = self.nrows // 6
        val = np.zeros(ncell); ind = np.zeros(ncell, np.int32); ln = C.c_int(0)
        getattr(self.L, "__m_thcm_utils_MOD_intcond_scaling")(val.ctypes.data_as(C.POINTER(C.c_double)), ind.ctypes.data_as(C.POINTER(C.c_int)), C.byref(ln))
        return val[:ln.value], ind[:ln.value]

    def recompute_scaling(self):
        """THCM::RecomputeScaling on one process (THCM.C:1694-1750) through the m_scaling symbols"""
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        db = np.zeros(36)
        getattr(self.L, "__m_scaling_MOD_average_block")(dp(db))
        rowscal, colscal = np.full(self.nrows, np.nan), np.full(self.nrows, np.nan)
        getattr(self.L, "__m_scaling_MOD_compute")(dp(db), dp(rowscal), dp(colscal))
        r, c = 1.0 / rowscal, 1.0 / colscal
        for v in (r, c):
            mean = 0.5 * (v[4::6] + v[5::6])
            v[4::6] = mean
            v[5::6] = mean
        return db.reshape(6, 6).T.copy(), r, c

    # ---- row N2: m_inserts, set_*_parameters_, getdeps_, m_probe (Ocean::synchronize / getBlock, Ocean.C:1443-1746) ----
    INSERTS = dict(tatm="atmosphere_t", qatm="atmosphere_q", albe="atmosphere_a", patm="atmosphere_p", qsa="seaice_q", msi="seaice_m", gsi="seaice_g")

    def insert(self, which, field):
        f = np.ascontiguousarray(field, np.float64).ravel()
        getattr(self.L, "__m_inserts_MOD_insert_" + self.INSERTS[which])(f.ctypes.data_as(C.POINTER(C.c_double)))

    def set_atmos_parameters(self, pars):
        from oracle_lib import Oracle
        a = np.array([float(pars[k]) for k in Oracle.ATMOS_PARS])
        self.L.set_atmos_parameters_(a.ctypes.data_as(C.c_void_p))

    def set_seaice_parameters(self, pars):
        from oracle_lib import Oracle
        a = np.array([float(pars[k]) for k in Oracle.SEAICE_PARS])
        self.L.set_seaice_parameters_(a.ctypes.data_as(C.c_void_p))

    def getdeps(self):
        v = [C.c_double(0.0) for _ in range(7)]
        self.L.getdeps_(*[C.byref(q) for q in v])
        return dict(zip(("Ooa", "Os", "nus", "eta", "lvsc", "qdim", "pQSnd"), (q.value for q in v)))

    def get_derivatives(self, un):
        un = np.ascontiguousarray(un, np.float64)
        nm = self.cfg["n"] * self.cfg["m"]
        out = [np.full(nm, np.nan) for _ in range(4)]
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        getattr(self.L, "__m_probe_MOD_get_derivatives")(dp(un), *[dp(o) for o in out])
        return out

    def close(self):
        self.L.finalize_()
        getattr(self.L, "__m_global_MOD_finalize")()
