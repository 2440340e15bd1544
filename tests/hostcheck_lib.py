"""ctypes binding of tests/host_rowcheck.cpp: the product's __host__ __device__ row code compiled for the CPU.

Test harness only -- lets the CPU test-suite check the arithmetic the CUDA kernels run against the oracle.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_lib import Config, PAR, ROOT

_LIB = None
_SO = os.path.join(ROOT, "tests", "_build", "libhostcheck.so")
_SRCS = [os.path.join(ROOT, "tests", "host_rowcheck.cpp"), os.path.join(ROOT, "iemic_b200", "csrc", "thcm_host.cpp")]
_DEPS = _SRCS + [os.path.join(ROOT, "iemic_b200", "csrc", f) for f in
                 ("thcm_rows.cuh", "thcm_stage.cuh", "thcm_cell.cuh", "thcm_coupled.cuh", "thcm_host.h")]


def lib():
    global _LIB
    if _LIB is None:
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        if (not os.path.exists(_SO)) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in _DEPS):
            subprocess.run(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-fPIC", "-shared", "-o", _SO] + _SRCS,
                           check=True)
        L = C.CDLL(_SO)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.hostcheck_create.restype = C.c_void_p
        L.hostcheck_create.argtypes = [C.POINTER(Config), ip, C.c_char_p, C.c_int]
        L.hostcheck_create_dist.restype = C.c_void_p
        L.hostcheck_create_dist.argtypes = [C.POINTER(Config), ip, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.hostcheck_nrows.argtypes = [C.c_void_p]
        L.hostcheck_nrows.restype = C.c_longlong
        L.hostcheck_nvec.argtypes = [C.c_void_p]
        L.hostcheck_nvec.restype = C.c_longlong
        L.hostcheck_gid.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
        L.hostcheck_block.argtypes = [C.c_void_p, ip]
        L.hostcheck_halo_npeers.argtypes = [C.c_void_p]
        L.hostcheck_halo_peer.argtypes = [C.c_void_p, C.c_int, ip]
        L.hostcheck_halo_lists.argtypes = [C.c_void_p, C.c_int, ip, ip]
        L.hostcheck_destroy.argtypes = [C.c_void_p]
        L.hostcheck_set_par.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_char_p, C.c_int]
        L.hostcheck_get_par.argtypes = [C.c_void_p, C.c_int]
        L.hostcheck_get_par.restype = C.c_double
        L.hostcheck_nnz.argtypes = [C.c_void_p]
        L.hostcheck_nnz.restype = C.c_longlong
        L.hostcheck_rowintcon.argtypes = [C.c_void_p]
        L.hostcheck_graph.argtypes = [C.c_void_p, ip, ip]
        L.hostcheck_frc.argtypes = [C.c_void_p, dp]
        L.hostcheck_sig.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong)]
        L.hostcheck_intcond.argtypes = [C.c_void_p, dp]
        L.hostcheck_jac.argtypes = [C.c_void_p, dp, dp, dp]
        L.hostcheck_rhs.argtypes = [C.c_void_p, dp, dp]
        L.hostcheck_scaling.argtypes = [C.c_void_p, dp, dp, dp, dp]
        L.hostcheck_scal.argtypes = [dp, dp, dp]
        L.hostcheck_insert.argtypes = [C.c_void_p, C.c_int, dp, C.c_char_p, C.c_int]
        L.hostcheck_set_atmos_parameters.argtypes = [C.c_void_p, dp]
        L.hostcheck_set_seaice_parameters.argtypes = [C.c_void_p, dp]
        L.hostcheck_getdeps.argtypes = [C.c_void_p, dp]
        L.hostcheck_suno.argtypes = [C.c_void_p, dp]
        L.hostcheck_dftdm.argtypes = [C.c_void_p, dp, dp]
        L.hostcheck_atmos_block.argtypes = [C.c_void_p, ip, ip, ip, dp]
        L.hostcheck_seaice_block.argtypes = [C.c_void_p, dp, ip, ip, ip, dp]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


class HostCheck:
    def __init__(self, cfg, landm, rank=0, nranks=1):
        self.c = Config(**cfg)
        self.landm = np.ascontiguousarray(landm, dtype=np.int32).ravel()
        err = C.create_string_buffer(512)
        self.h = lib().hostcheck_create_dist(C.byref(self.c), _ip(self.landm), rank, nranks, err, 512)
        if not self.h:
            raise RuntimeError(err.value.decode())
        self.ndim = lib().hostcheck_nrows(self.h)       # local rows
        self.nvec = lib().hostcheck_nvec(self.h)        # stored vector length (owned + halo ring)
        self._graph = None

    def gid(self):
        g = np.zeros(self.nvec, np.int64)
        lib().hostcheck_gid(self.h, g.ctypes.data_as(C.POINTER(C.c_longlong)))
        return g

    def halo_plan(self):
        """[(peer rank, send cell indices, recv cell indices)]"""
        out = []
        for q in range(lib().hostcheck_halo_npeers(self.h)):
            info = np.zeros(3, np.int32)
            lib().hostcheck_halo_peer(self.h, q, _ip(info))
            send, recv = np.zeros(info[1], np.int32), np.zeros(info[2], np.int32)
            lib().hostcheck_halo_lists(self.h, q, _ip(send), _ip(recv))
            out.append((int(info[0]), send, recv))
        return out

    def block(self):
        b = np.zeros(6, np.int32)
        lib().hostcheck_block(self.h, _ip(b))
        return tuple(int(v) for v in b)

    def __del__(self):
        if getattr(self, "h", None):
            lib().hostcheck_destroy(self.h)
            self.h = None

    def set_par(self, name, value):
        err = C.create_string_buffer(512)
        if lib().hostcheck_set_par(self.h, PAR[name] if isinstance(name, str) else name, float(value), err, 512) != 0:
            raise RuntimeError(err.value.decode())

    def get_par(self, name):
        return lib().hostcheck_get_par(self.h, PAR[name] if isinstance(name, str) else name)

    # coupling: same names and orders as oracle_lib.Oracle
    def insert(self, which, field):
        from oracle_lib import Oracle
        f = np.ascontiguousarray(field, np.float64).ravel()
        err = C.create_string_buffer(512)
        if lib().hostcheck_insert(self.h, Oracle.FIELDS[which], _dp(f), err, 512) != 0:
            raise RuntimeError(err.value.decode())

    def set_atmos_parameters(self, pars):
        from oracle_lib import Oracle
        a = np.array([float(pars[k]) for k in Oracle.ATMOS_PARS])
        lib().hostcheck_set_atmos_parameters(self.h, _dp(a))

    def set_seaice_parameters(self, pars):
        from oracle_lib import Oracle
        a = np.array([float(pars[k]) for k in Oracle.SEAICE_PARS])
        lib().hostcheck_set_seaice_parameters(self.h, _dp(a))

    def getdeps(self):
        out = np.zeros(7)
        lib().hostcheck_getdeps(self.h, _dp(out))
        return dict(zip(("Ooa", "Os", "nus", "eta", "lvsc", "qdim", "pQSnd"), out))

    def suno(self):
        out = np.zeros(self.c.m)
        lib().hostcheck_suno(self.h, _dp(out))
        return out

    def dftdm(self, x):
        x = np.ascontiguousarray(x, np.float64)
        i0, j0, nloc, mloc, _, _ = self.block()
        out = np.zeros(nloc * mloc)
        lib().hostcheck_dftdm(self.h, _dp(x), _dp(out))
        return out

    def _block(self, fn, *lead):
        nm = self.c.n * self.c.m
        rows, beg, jco, co = np.zeros(nm, np.int32), np.zeros(nm + 1, np.int32), np.zeros(3 * nm, np.int32), np.zeros(3 * nm)
        k = fn(self.h, *lead, _ip(rows), _ip(beg), _ip(jco), _dp(co))
        return rows[:k].copy(), beg[:k + 1].copy(), jco[:beg[k]].copy(), co[:beg[k]].copy()

    def atmos_block(self):
        return self._block(lib().hostcheck_atmos_block)

    def seaice_block(self, dftdm):
        d = np.ascontiguousarray(dftdm, np.float64)
        return self._block(lib().hostcheck_seaice_block, _dp(d))

    def graph(self):
        if self._graph is None:
            nnz = lib().hostcheck_nnz(self.h)
            rp = np.zeros(self.ndim + 1, np.int32)
            col = np.zeros(nnz, np.int32)
            lib().hostcheck_graph(self.h, _ip(rp), _ip(col))
            self._graph = (rp, col)
        return self._graph

    def jac(self, x):
        x = np.ascontiguousarray(x, np.float64)
        rp, col = self.graph()
        val = np.zeros(col.size)
        diagB = np.zeros(self.ndim)
        lib().hostcheck_jac(self.h, _dp(x), _dp(val), _dp(diagB))
        return val, diagB

    def rhs(self, x):
        x = np.ascontiguousarray(x, np.float64)
        F = np.zeros(self.ndim)
        lib().hostcheck_rhs(self.h, _dp(x), _dp(F))
        return F

    def scaling(self, x):
        """(db[ii, jj], rs6, cs6) of the product's scaling path evaluated on the CPU; None if the block is singular"""
        x = np.ascontiguousarray(x, np.float64)
        db, rs, cs = np.zeros(36), np.zeros(6), np.zeros(6)
        ok = lib().hostcheck_scaling(self.h, _dp(x), _dp(db), _dp(rs), _dp(cs))
        return (db.reshape(6, 6).T.copy(), rs, cs) if ok else None

    def frc(self):
        f = np.zeros(self.ndim)
        lib().hostcheck_frc(self.h, _dp(f))
        return f

    def rowintcon(self):
        return lib().hostcheck_rowintcon(self.h)


def scal(db):
    """thcm::scal of the product on a 6x6 array db[ii, jj]: (dr, dc) or None if singular"""
    a = np.ascontiguousarray(np.asarray(db, np.float64).T).ravel()  # column-major
    dr, dc = np.zeros(6), np.zeros(6)
    return (dr, dc) if lib().hostcheck_scal(_dp(a), _dp(dr), _dp(dc)) else None
