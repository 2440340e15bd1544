"""Builds libiemic_b200.so (the C-ABI shared library of include/iemic_b200.h) in-tree with nvcc for sm_100a.

No torch involvement: plain `nvcc -shared`.  The library lands in iemic_b200/lib/ so that it travels to the GPU box
with the repository snapshot.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libiemic_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
# (source, extra flags).  thcm_kernels.cu: no FMA contraction -> bit parity with the FMA-free reference arithmetic.
UNITS = [
    ("thcm_kernels.cu", ["-fmad=false"]),
    ("spmv.cu", []),
    ("spmv_tile.cu", ["-fmad=false"]),
    ("gmres.cu", []),
    ("gmres_cycle.cu", []),
    ("gmres_cluster.cu", []),
    ("scaling.cu", ["-fmad=false"]),
    ("coupled.cu", ["-fmad=false"]),
    ("dist.cu", []),
    ("capi.cu", ["-Xcompiler", "-fvisibility=default"]),
    # host-side table construction: host compiler must not contract a*b+c either
    ("thcm_host.cpp", ["-Xcompiler", "-ffp-contract=off"]),
]
HEADERS = ["thcm_rows.cuh", "thcm_stage.cuh", "thcm_cell.cuh", "thcm_host.h", "kernels.h", "dist.h", "peer.cuh", "gmres_givens.cuh", "gmres_guard.h",
           os.path.join("..", "..", "include", "iemic_b200.h")]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _digest(paths, flags):
    h = hashlib.sha256()
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def build(force=False, verbose=False, defines=(), tag=None):
    """tag/defines: build an experimental variant into lib/variants/libiemic_b200_<tag>.so (tools/ only)"""
    global LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(HERE, "build" if not tag else os.path.join("build", tag))
    lib = LIB if not tag else os.path.join(LIBDIR, "variants", "libiemic_b200_%s.so" % tag)
    os.makedirs(os.path.dirname(lib), exist_ok=True)
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    objs = []
    rebuilt = False
    procs = []
    for src, extra in UNITS:
        sp = os.path.join(CSRC, src)
        obj = os.path.join(objdir, src + ".o")
        stamp = obj + ".sha"
        flags = ARCH + COMMON + extra + ["-D" + d for d in defines]
        dig = _digest([sp] + hdrs, flags)
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
            continue
        cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", sp, "-o", obj]
        procs.append((subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True), stamp, dig, src))
        rebuilt = True
    for p, stamp, dig, src in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on " + src)
        if verbose:
            sys.stderr.write(out)
        with open(stamp, "w") as f:
            f.write(dig)
    if rebuilt or not os.path.exists(lib):
        cmd = [_nvcc()] + ARCH + ["-shared", "-o", lib] + objs + ["-cudart", "static", "-ldl"]
        subprocess.run(cmd, check=True)
    if not tag:
        build_fortran_abi(lib, force or rebuilt)
    return lib


FORTRAN_LIB = os.path.join(LIBDIR, "libthcm_b200_fortran.so")


def build_fortran_abi(core_lib, force=False):
    """Seam S1 (include/thcm_fortran_abi.h): the Fortran symbols THCM.C binds, as a thin host library over the C ABI.
    A separate .so: names like init_ / rhs_ / matrix_ do not belong in the namespace of every process that loads the
    core library."""
    src = os.path.join(CSRC, "fortran_abi.cpp")
    inc = [os.path.join(HERE, "..", "include", "thcm_fortran_abi.h"), os.path.join(HERE, "..", "include", "iemic_b200.h")]
    flags = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-shared"]
    dig = _digest([src] + inc, flags)
    stamp = os.path.join(HERE, "build", "fortran_abi.sha")
    if not force and os.path.exists(FORTRAN_LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return FORTRAN_LIB
    cmd = ["g++"] + flags + ["-o", FORTRAN_LIB, src, "-L" + LIBDIR, "-liemic_b200", "-Wl,-rpath,$ORIGIN", "-ldl"]
    subprocess.run(cmd, check=True)
    with open(stamp, "w") as f:
        f.write(dig)
    return FORTRAN_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
