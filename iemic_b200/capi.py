"""ctypes binding of libiemic_b200.so (include/iemic_b200.h).

This is the only way Python reaches the hot path: every call below crosses the C ABI.  There is no CPU
fallback -- if the library is missing or no CUDA device is present, creation fails loudly.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("IEMIC_B200_LIB") or os.path.join(HERE, "lib", "libiemic_b200.so")  # override: kernel-variant experiments
_LIB = None

# THCM::par2int (src/ocean/THCM.C:1754-1869 of the reference) / par.F90:37-66
PAR_NAMES = {
    "AL_T": 1, "Rayleigh-Number": 2, "Vertical Ekman-Number": 3, "Horizontal Ekman-Number": 4, "Rossby-Number": 5,
    "MIXP": 6, "RESC": 7, "SPL1": 8, "Salinity Homotopy": 9, "Solar Forcing": 10, "Horizontal Peclet-Number": 11,
    "Vertical Peclet-Number": 12, "P_VC": 13, "LAMB": 14, "Salinity Forcing": 15, "Wind Forcing": 16,
    "Temperature Forcing": 17, "Nonlinear Factor": 18, "Combined Forcing": 19, "ARCL": 20, "NLES": 21, "IFRICB": 22,
    "CONT": 23, "Energy": 24, "ALPC": 25, "CMPR": 26, "Flux Perturbation": 27, "Salinity Perturbation": 28, "MKAP": 29,
    "SPL2": 30,
}
PAR = dict(AL_T=1, RAYL=2, EK_V=3, EK_H=4, ROSB=5, MIXP=6, RESC=7, SPL1=8, HMTP=9, SUNP=10, PE_H=11,
           PE_V=12, P_VC=13, LAMB=14, SALT=15, WIND=16, TEMP=17, BIOT=18, COMB=19, ARCL=20, NLES=21,
           IFRICB=22, CONT=23, ENER=24, ALPC=25, CMPR=26, FPER=27, SPER=28, MKAP=29, SPL2=30)

MGS, CGS2, CGS2F = 0, 1, 2
PREC_NONE, PREC_BLOCKJACOBI, PREC_CALLBACK = 0, 1, 2
PRECOND_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p)


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


class B200Error(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error("libiemic_b200.so is not built (run `python -m iemic_b200.build`); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp, dp, ip, i64p = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int64)
        L.iemic_b200_last_error.restype = C.c_char_p
        L.iemic_b200_version.restype = C.c_char_p
        sig = {
            "thcm_b200_create": [C.POINTER(Config), ip, C.POINTER(vp)],
            "thcm_b200_create_dist": [C.POINTER(Config), ip, C.c_int, C.c_int, vp, C.POINTER(vp)],
            "thcm_b200_comm_unique_id": [vp],
            "thcm_b200_destroy": [vp], "thcm_b200_set_stream": [vp, vp], "thcm_b200_sync": [vp],
            "thcm_b200_set_par": [vp, C.c_int, C.c_double], "thcm_b200_get_par": [vp, C.c_int, dp],
            "thcm_b200_set_landmask": [vp, ip], "thcm_b200_set_intcond_correction": [vp, C.c_double],
            "thcm_b200_sizes": [vp, i64p, i64p, i64p, i64p],
            "thcm_b200_block": [vp, ip, ip, ip, ip, ip, ip],
            "thcm_b200_global_ids": [vp, i64p],
            "thcm_b200_rhs": [vp, vp, vp], "thcm_b200_jacobian": [vp, vp], "thcm_b200_halo_exchange": [vp, vp],
            "thcm_b200_csr": [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)],
            "thcm_b200_graph_host": [vp, ip, ip], "thcm_b200_values_host": [vp, dp, dp],
            "thcm_b200_set_values_host": [vp, dp], "thcm_b200_forcing_host": [vp, dp],
            "thcm_b200_intcond_host": [vp, dp, ip],
            "thcm_b200_rhs_host": [vp, vp, vp], "thcm_b200_jacobian_host": [vp, vp, vp, vp],
            "thcm_b200_assemble_host": [vp, vp, vp, C.c_int], "thcm_b200_assemble": [vp, vp, vp, C.c_int], "thcm_b200_exchange_mode": [vp], "thcm_b200_evaluate_b": [vp, vp],
            "thcm_b200_spmv": [vp, vp, vp], "thcm_b200_spmv_host": [vp, vp, vp],
            "vec_b200_dot": [vp, vp, vp, dp], "vec_b200_norm2": [vp, vp, dp],
            "vec_b200_dot_noalloc": [vp, vp, C.c_int64, vp, vp], "vec_b200_norminf": [vp, vp, dp],
            "thcm_b200_upload": [vp, vp, vp, C.c_int64], "thcm_b200_download": [vp, vp, vp, C.c_int64],
            "vec_b200_update": [vp, C.c_double, vp, C.c_double, vp], "vec_b200_scale": [vp, C.c_double, vp],
            "vec_b200_putscalar": [vp, C.c_double, vp],
            "vec_b200_multiply": [vp, C.c_double, vp, vp, C.c_double, vp], "vec_b200_reciprocal": [vp, vp, vp],
            "thcm_b200_average_block": [vp, dp], "thcm_b200_scal": [dp, dp, dp], "thcm_b200_recompute_scaling": [vp, vp, vp, dp, dp],
            "thcm_b200_left_scale": [vp, vp], "thcm_b200_block_jacobi": [vp, vp, ip],
            "gmres_b200_solve": [vp, vp, vp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, vp, C.c_int, dp, ip],
            "gmres_b200_solve_host": [vp, vp, vp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, vp, C.c_int, dp, ip],
            "gmres_b200_set_precond": [vp, PRECOND_FN, vp], "vec_b200_block_jacobi_apply": [vp, vp, vp, vp],
            "thcm_b200_launch_count": [vp, i64p], "thcm_b200_timing_enable": [vp, C.c_int],
            "thcm_b200_timing_reset": [vp], "thcm_b200_timing_get": [vp, C.c_int, dp, i64p],
            "iemic_b200_malloc": [C.POINTER(vp), C.c_int64], "iemic_b200_free": [vp],
            "iemic_b200_memcpy_h2d": [vp, vp, C.c_int64], "iemic_b200_memcpy_d2h": [vp, vp, C.c_int64],
            "iemic_b200_device_count": [],
            # row N2: coupling to an atmosphere / sea-ice model
            "thcm_b200_insert_field": [vp, C.c_int, dp], "thcm_b200_set_atmos_parameters": [vp, dp], "thcm_b200_set_seaice_parameters": [vp, dp],
            "thcm_b200_getdeps": [vp, dp], "thcm_b200_get_suno": [vp, dp], "thcm_b200_get_dftdm": [vp, vp, dp],
            "thcm_b200_atmos_block": [vp, ip, ip, ip, dp, ip], "thcm_b200_seaice_block": [vp, vp, ip, ip, ip, dp, ip],
            "coupling_b200_create": [vp, C.c_int, ip, ip, ip, dp, C.c_int64, C.POINTER(vp)], "coupling_b200_apply": [vp, vp, vp, C.c_int],
        }
        for name, args in sig.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        L.coupling_b200_destroy.argtypes = [vp]
        L.coupling_b200_destroy.restype = None
        _LIB = L
    return _LIB


EXPORTED = ["iemic_b200_last_error", "iemic_b200_version", "iemic_b200_device_count",
            "thcm_b200_create", "thcm_b200_create_dist", "thcm_b200_comm_unique_id", "thcm_b200_destroy",
            "thcm_b200_set_stream", "thcm_b200_sync", "thcm_b200_set_par", "thcm_b200_get_par", "thcm_b200_set_landmask",
            "thcm_b200_set_intcond_correction", "thcm_b200_sizes", "thcm_b200_block", "thcm_b200_global_ids",
            "thcm_b200_rhs", "thcm_b200_jacobian", "thcm_b200_halo_exchange", "thcm_b200_csr", "thcm_b200_graph_host",
            "thcm_b200_values_host", "thcm_b200_set_values_host", "thcm_b200_forcing_host", "thcm_b200_intcond_host",
            "thcm_b200_rhs_host", "thcm_b200_jacobian_host", "thcm_b200_assemble_host", "thcm_b200_assemble", "thcm_b200_exchange_mode", "thcm_b200_evaluate_b", "thcm_b200_spmv", "thcm_b200_spmv_host",
            "vec_b200_dot", "vec_b200_dot_noalloc", "vec_b200_norm2", "vec_b200_norminf", "thcm_b200_upload", "thcm_b200_download", "vec_b200_update", "vec_b200_scale", "vec_b200_putscalar",
            "vec_b200_multiply", "vec_b200_reciprocal", "thcm_b200_average_block", "thcm_b200_scal", "thcm_b200_recompute_scaling",
            "thcm_b200_left_scale", "thcm_b200_block_jacobi",
            "gmres_b200_solve", "gmres_b200_solve_host", "gmres_b200_set_precond", "vec_b200_block_jacobi_apply", "thcm_b200_launch_count", "thcm_b200_timing_enable",
            "thcm_b200_timing_reset", "thcm_b200_timing_get", "iemic_b200_malloc", "iemic_b200_free",
            "iemic_b200_memcpy_h2d", "iemic_b200_memcpy_d2h",
            "thcm_b200_insert_field", "thcm_b200_set_atmos_parameters", "thcm_b200_set_seaice_parameters", "thcm_b200_getdeps", "thcm_b200_get_suno",
            "thcm_b200_get_dftdm", "thcm_b200_atmos_block", "thcm_b200_seaice_block", "coupling_b200_create", "coupling_b200_apply", "coupling_b200_destroy"]


def check(rc):
    if rc < 0:
        raise B200Error(lib().iemic_b200_last_error().decode() or ("error %d" % rc))
    return rc


def _hp(a):
    """host pointer of a contiguous numpy array"""
    return C.c_void_p(a.ctypes.data)


def _ptr(t):
    """device pointer of a torch tensor / raw int"""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(int(t))


class Ocean:
    """Host-side mirror of the hot members of the reference's Ocean model (src/ocean/Ocean.H): computeRHS,
    computeJacobian, applyMatrix, solve, getPar/setPar -- backed by one thcm_b200 handle."""

    def __init__(self, cfg, landm, rank=0, nranks=1, nccl_id=None):
        L = lib()
        self.cfg = dict(cfg)
        self._c = Config(**cfg)
        lm = np.ascontiguousarray(landm, dtype=np.int32).ravel()
        if lm.size != (cfg["n"] + 2) * (cfg["m"] + 2) * (cfg["l"] + 2):
            raise ValueError("land mask must have (n+2)(m+2)(l+2) entries")
        h = C.c_void_p()
        if nranks == 1:
            check(L.thcm_b200_create(C.byref(self._c), lm.ctypes.data_as(C.POINTER(C.c_int)), C.byref(h)))
        else:
            idp = C.c_void_p(nccl_id.ctypes.data) if nccl_id is not None else None
            check(L.thcm_b200_create_dist(C.byref(self._c), lm.ctypes.data_as(C.POINTER(C.c_int)), rank, nranks, idp, C.byref(h)))
        self.h = h
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(L.thcm_b200_sizes(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        self.nrows, self.nvec, self.nnz, self.nglobal = a.value, b.value, c.value, d.value
        self.ncell = self.nrows // 6

    @classmethod
    def from_xml(cls, ocean_params_file, mask_dir=".", **kw):
        """Ocean::Ocean(comm, oceanParams) for this path: the reference's ocean parameter file and land-mask file, unchanged
        (iemic_b200.params.ocean_setup); "Starting Parameters" are applied."""
        from . import params
        cfg, mask, start = params.ocean_setup(ocean_params_file, mask_dir)
        o = cls(cfg, mask, **kw)
        for name, value in start.items():
            o.setPar(name, value)
        return o

    def close(self):
        if getattr(self, "h", None):
            lib().thcm_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters (Ocean::setPar/getPar -> THCM::setParameter) ----
    def setPar(self, name, value):
        idx = PAR[name] if name in PAR else (PAR_NAMES[name] if isinstance(name, str) else int(name))
        check(lib().thcm_b200_set_par(self.h, idx, float(value)))

    def getPar(self, name):
        idx = PAR[name] if name in PAR else (PAR_NAMES[name] if isinstance(name, str) else int(name))
        v = C.c_double()
        check(lib().thcm_b200_get_par(self.h, idx, C.byref(v)))
        return v.value

    # ---- row N2: Ocean::synchronize(atmos / seaice), Ocean::getBlock, THCM::getDerivatives ----
    FIELDS = dict(tatm=0, qatm=1, albe=2, patm=3, qsa=4, msi=5, gsi=6)
    ATMOS_PARS = ("tdim", "qdim", "nuq", "eta", "dqso", "dqsi", "dqdt", "Eo0", "Ei0", "Cs", "t0o", "t0i", "a0", "da", "tauf", "tauc", "comb", "albf")
    SEAICE_PARS = ("zeta", "a0", "Lf", "s0", "rhoo", "qvar", "q0")

    def insert(self, which, field):
        f = np.ascontiguousarray(field, np.float64).ravel()
        assert f.size == self.cfg["n"] * self.cfg["m"]
        check(lib().thcm_b200_insert_field(self.h, self.FIELDS[which], f.ctypes.data_as(C.POINTER(C.c_double))))

    def set_atmos_parameters(self, pars):
        a = np.array([float(pars[k]) for k in self.ATMOS_PARS])
        check(lib().thcm_b200_set_atmos_parameters(self.h, a.ctypes.data_as(C.POINTER(C.c_double))))

    def set_seaice_parameters(self, pars):
        a = np.array([float(pars[k]) for k in self.SEAICE_PARS])
        check(lib().thcm_b200_set_seaice_parameters(self.h, a.ctypes.data_as(C.POINTER(C.c_double))))

    def getdeps(self):
        out = np.zeros(7)
        check(lib().thcm_b200_getdeps(self.h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return dict(zip(("Ooa", "Os", "nus", "eta", "lvsc", "qdim", "pQSnd"), out))

    def suno(self):
        out = np.zeros(self.cfg["n"] * self.cfg["m"])
        check(lib().thcm_b200_get_suno(self.h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def dftdm(self, d_x):
        i0, j0, nloc, mloc, _, _ = self.block()
        out = np.zeros(nloc * mloc)
        check(lib().thcm_b200_get_dftdm(self.h, _ptr(d_x), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def _block(self, fn, *lead):
        nm = self.cfg["n"] * self.cfg["m"]
        rows, beg, jco, co = np.zeros(nm, np.int32), np.zeros(nm + 1, np.int32), np.zeros(3 * nm, np.int32), np.zeros(3 * nm)
        nr = C.c_int(0)
        ip, dp = C.POINTER(C.c_int), C.POINTER(C.c_double)
        check(fn(self.h, *lead, rows.ctypes.data_as(ip), beg.ctypes.data_as(ip), jco.ctypes.data_as(ip), co.ctypes.data_as(dp), C.byref(nr)))
        k = nr.value
        return rows[:k].copy(), beg[:k + 1].copy(), jco[:beg[k]].copy(), co[:beg[k]].copy()

    def atmos_block(self):
        """Ocean::getBlock(Atmosphere), compact: (rows, beg, jco, co)"""
        return self._block(lib().thcm_b200_atmos_block)

    def seaice_block(self, d_x):
        return self._block(lib().thcm_b200_seaice_block, _ptr(d_x))

    def coupling_block(self, rows, beg, jco, co, ncols):
        return CouplingBlock(self, rows, beg, jco, co, ncols)

    def set_landmask(self, landm):
        lm = np.ascontiguousarray(landm, dtype=np.int32).ravel()
        check(lib().thcm_b200_set_landmask(self.h, lm.ctypes.data_as(C.POINTER(C.c_int))))
        a, b, c, d = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(lib().thcm_b200_sizes(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        self.nrows, self.nvec, self.nnz, self.nglobal = a.value, b.value, c.value, d.value

    def set_intcond_correction(self, corr):
        check(lib().thcm_b200_set_intcond_correction(self.h, float(corr)))

    # ---- host-buffer entry points (numpy in / numpy out) ----
    def graph(self):
        rp = np.zeros(self.nrows + 1, np.int32)
        col = np.zeros(self.nnz, np.int32)
        check(lib().thcm_b200_graph_host(self.h, rp.ctypes.data_as(C.POINTER(C.c_int)), col.ctypes.data_as(C.POINTER(C.c_int))))
        return rp, col

    def global_ids(self):
        g = np.zeros(self.nvec, np.int64)
        check(lib().thcm_b200_global_ids(self.h, g.ctypes.data_as(C.POINTER(C.c_int64))))
        return g

    def block(self):
        v = [C.c_int() for _ in range(6)]
        check(lib().thcm_b200_block(self.h, *[C.byref(q) for q in v]))
        return tuple(q.value for q in v)

    def computeRHS_host(self, x, out=None):
        x = np.ascontiguousarray(x, np.float64)
        F = np.empty(self.nrows) if out is None else out
        check(lib().thcm_b200_rhs_host(self.h, _hp(x), _hp(F)))
        return F

    def computeJacobian_host(self, x, want_values=True):
        x = np.ascontiguousarray(x, np.float64)
        if not want_values:
            check(lib().thcm_b200_jacobian_host(self.h, _hp(x), None, None))
            return None
        val = np.empty(self.nnz)
        diagB = np.empty(self.nrows)
        check(lib().thcm_b200_jacobian_host(self.h, _hp(x), _hp(val), _hp(diagB)))
        return val, diagB

    def assemble_host(self, x, F=None, jacobian=True):
        """computeRHS + computeJacobian on one state: one upload, F downloaded under the Jacobian kernel"""
        check(lib().thcm_b200_assemble_host(self.h, _hp(x), _hp(F) if F is not None else None, int(jacobian)))
        return F

    def values_host(self):
        val = np.empty(self.nnz)
        diagB = np.empty(self.nrows)
        check(lib().thcm_b200_values_host(self.h, val.ctypes.data_as(C.POINTER(C.c_double)), diagB.ctypes.data_as(C.POINTER(C.c_double))))
        return val, diagB

    def set_values_host(self, val):
        val = np.ascontiguousarray(val, np.float64)
        assert val.size == self.nnz
        check(lib().thcm_b200_set_values_host(self.h, val.ctypes.data_as(C.POINTER(C.c_double))))

    def forcing_host(self):
        f = np.empty(self.nrows)
        check(lib().thcm_b200_forcing_host(self.h, f.ctypes.data_as(C.POINTER(C.c_double))))
        return f

    def applyMatrix_host(self, x):
        x = np.ascontiguousarray(x, np.float64)
        y = np.empty(self.nrows)
        check(lib().thcm_b200_spmv_host(self.h, _hp(x), _hp(y)))
        return y

    def solve_host(self, b, x0=None, m=30, maxit=30, tol=1e-8, ortho=CGS2, prec=PREC_NONE, blockinv=None, check_every=1):
        b = np.ascontiguousarray(b, np.float64)
        x = np.zeros(self.nrows) if x0 is None else np.array(x0, np.float64)
        hist = np.zeros(maxit + m + 4)
        it = C.c_int(0)
        bi = None
        if blockinv is not None:
            blockinv = np.ascontiguousarray(blockinv, np.float64)
            bi = _hp(blockinv)
        rc = check(lib().gmres_b200_solve_host(self.h, _hp(b), _hp(x), m, maxit, tol, ortho, prec, bi, check_every,
                                               hist.ctypes.data_as(C.POINTER(C.c_double)), C.byref(it)))
        return rc, x, hist, it.value

    # ---- device entry points (torch tensors or raw device pointers) ----
    def computeRHS(self, d_x, d_F):
        check(lib().thcm_b200_rhs(self.h, _ptr(d_x), _ptr(d_F)))

    def computeJacobian(self, d_x):
        check(lib().thcm_b200_jacobian(self.h, _ptr(d_x)))

    def evaluateB(self, d_diagB=None):
        check(lib().thcm_b200_evaluate_b(self.h, _ptr(d_diagB)))

    def exchange_mode(self):
        """0 single GPU, 1 NCCL, 2 peer memory over NVLink"""
        return int(lib().thcm_b200_exchange_mode(self.h))

    def assemble(self, d_x, d_F=None, jacobian=True):
        """computeRHS and/or computeJacobian on one device-resident state, halo exchange overlapped"""
        check(lib().thcm_b200_assemble(self.h, _ptr(d_x), _ptr(d_F), int(jacobian)))

    def applyMatrix(self, d_x, d_y):
        check(lib().thcm_b200_spmv(self.h, _ptr(d_x), _ptr(d_y)))

    def halo_exchange(self, d_x):
        check(lib().thcm_b200_halo_exchange(self.h, _ptr(d_x)))

    def solve(self, d_b, d_x, m=30, maxit=30, tol=1e-8, ortho=CGS2, prec=PREC_NONE, d_blockinv=None, check_every=1):
        hist = np.zeros(maxit + m + 4)
        it = C.c_int(0)
        rc = check(lib().gmres_b200_solve(self.h, _ptr(d_b), _ptr(d_x), m, maxit, tol, ortho, prec, _ptr(d_blockinv), check_every,
                                          hist.ctypes.data_as(C.POINTER(C.c_double)), C.byref(it)))
        return rc, hist, it.value

    def set_precond(self, fn):
        """fn(d_v_ptr, d_z_ptr, n, stream_ptr) -> int: the right preconditioner z = M^-1 v used with prec=PREC_CALLBACK (None: remove)"""
        self._precond_cb = PRECOND_FN(lambda user, v, z, n, st: int(fn(v, z, n, st) or 0)) if fn is not None else C.cast(None, PRECOND_FN)
        check(lib().gmres_b200_set_precond(self.h, self._precond_cb, None))

    def dot(self, d_a, d_b):
        r = C.c_double()
        check(lib().vec_b200_dot(self.h, _ptr(d_a), _ptr(d_b), C.byref(r)))
        return r.value

    def norm2(self, d_a):
        r = C.c_double()
        check(lib().vec_b200_norm2(self.h, _ptr(d_a), C.byref(r)))
        return r.value

    def update(self, a, d_A, b, d_this):
        check(lib().vec_b200_update(self.h, float(a), _ptr(d_A), float(b), _ptr(d_this)))

    def multiply(self, a, d_A, d_B, b, d_this):
        check(lib().vec_b200_multiply(self.h, float(a), _ptr(d_A), _ptr(d_B), float(b), _ptr(d_this)))

    def reciprocal(self, d_A, d_this):
        check(lib().vec_b200_reciprocal(self.h, _ptr(d_A), _ptr(d_this)))

    # ---- scaling and point-block Jacobi on the current Jacobian (THCM::RecomputeScaling, Ocean::scaleProblem) ----
    def average_block(self):
        """m_scaling::average_block: 6x6 array db[ii, jj]"""
        db = np.zeros(36)
        check(lib().thcm_b200_average_block(self.h, db.ctypes.data_as(C.POINTER(C.c_double))))
        return db.reshape(6, 6).T.copy()  # column-major on the C side

    def recompute_scaling(self, d_rowscal=None, d_colscal=None):
        """THCM::RecomputeScaling; returns the six row and six column factors of ocean cells"""
        rs, cs = np.zeros(6), np.zeros(6)
        check(lib().thcm_b200_recompute_scaling(self.h, _ptr(d_rowscal), _ptr(d_colscal), rs.ctypes.data_as(C.POINTER(C.c_double)),
                                                cs.ctypes.data_as(C.POINTER(C.c_double))))
        return rs, cs

    def left_scale(self, d_scal):
        check(lib().thcm_b200_left_scale(self.h, _ptr(d_scal)))

    def block_jacobi(self, d_blockinv):
        ns = C.c_int(0)
        check(lib().thcm_b200_block_jacobi(self.h, _ptr(d_blockinv), C.byref(ns)))
        return ns.value

    def set_stream(self, stream_ptr):
        check(lib().thcm_b200_set_stream(self.h, C.c_void_p(int(stream_ptr))))

    def sync(self):
        check(lib().thcm_b200_sync(self.h))

    def launch_count(self):
        v = C.c_int64()
        check(lib().thcm_b200_launch_count(self.h, C.byref(v)))
        return v.value

    def timing(self, on):
        check(lib().thcm_b200_timing_enable(self.h, int(on)))

    def timing_reset(self):
        check(lib().thcm_b200_timing_reset(self.h))

    def timing_get(self, which):
        t, c = C.c_double(), C.c_int64()
        check(lib().thcm_b200_timing_get(self.h, which, C.byref(t), C.byref(c)))
        return t.value, c.value


class CouplingBlock:
    """CouplingBlock<Ocean, Model> on the device (src/coupledmodel/CouplingBlock.H)"""

    def __init__(self, ocean, rows, beg, jco, co, ncols):
        self.ocean = ocean  # keeps the handle alive
        rows, beg, jco = (np.ascontiguousarray(a, np.int32) for a in (rows, beg, jco))
        co = np.ascontiguousarray(co, np.float64)
        ip, dp = C.POINTER(C.c_int), C.POINTER(C.c_double)
        self.b = C.c_void_p()
        check(lib().coupling_b200_create(ocean.h, int(rows.size), rows.ctypes.data_as(ip), beg.ctypes.data_as(ip), jco.ctypes.data_as(ip),
                                         co.ctypes.data_as(dp), int(ncols), C.byref(self.b)))

    def apply(self, d_in, d_out, accumulate=False):
        check(lib().coupling_b200_apply(self.b, _ptr(d_in), _ptr(d_out), 1 if accumulate else 0))

    def __del__(self):
        if getattr(self, "b", None):
            lib().coupling_b200_destroy(self.b)
            self.b = None
