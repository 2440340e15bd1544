"""Builds tests/continuation_dropin.cpp: the reference's UNMODIFIED Continuation<Model> (read in place from the reference
tree, never copied) instantiated on the B200 host mirror.  The reference tree only exists in the build container, so the
binary goes to tests/_build/ (git-ignored, shipped to the GPU box with the snapshot) and the GPU test only runs it."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CONT = "/root/reference/src/continuation"
EXE = os.path.join(ROOT, "tests", "_build", "continuation_dropin")


def _fresh(exe, *deps):
    """exe exists and is newer than every file / every file of every directory in deps"""
    if not os.path.exists(exe):
        return False
    t = os.path.getmtime(exe)
    for d in deps:
        files = [os.path.join(d, f) for f in os.listdir(d)] if os.path.isdir(d) else [d]
        if any(os.path.isfile(f) and os.path.getmtime(f) > t for f in files):
            return False
    return True


def build_continuation_dropin():
    """-> path of the binary, or None when the reference tree is absent"""
    if not os.path.isdir(REF_CONT):
        return EXE if os.path.exists(EXE) else None
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    host = os.path.join(ROOT, "iemic_b200", "host")
    if _fresh(EXE, os.path.join(ROOT, "tests", "continuation_dropin.cpp"), os.path.join(ROOT, "tests", "continuation_cases.H"),
              os.path.join(ROOT, "tests", "compat"), host, os.path.join(ROOT, "include"), REF_CONT,
              os.path.join(ROOT, "iemic_b200", "lib", "libiemic_b200.so")):
        return EXE
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "tests", "compat"), "-I", REF_CONT,
           "-I", os.path.join(ROOT, "include"), "-I", host, os.path.join(ROOT, "tests", "continuation_dropin.cpp"),
           "-L", os.path.join(ROOT, "iemic_b200", "lib"), "-liemic_b200", "-ldl", "-Wl,-rpath,$ORIGIN/../../iemic_b200/lib", "-o", EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("continuation_dropin.cpp does not compile against the reference's Continuation.H:\n" + r.stderr[-4000:])
    return EXE


# reference code compiled here (Continuation.H / GMRESSolver.H, read in place) that does not touch the product: oracle/_ref/
# (git-ignored, shipped to the GPU box like the other built files)
ORACLE_EXE = os.path.join(ROOT, "oracle", "_ref", "continuation_oracle")


def build_continuation_oracle():
    """tests/continuation_oracle.cpp: the same unmodified Continuation<Model> on the CPU oracle (test infrastructure).
    -> path of the binary, or None when the reference tree is absent"""
    if not os.path.isdir(REF_CONT):
        return ORACLE_EXE if os.path.exists(ORACLE_EXE) else None
    os.makedirs(os.path.dirname(ORACLE_EXE), exist_ok=True)
    if _fresh(ORACLE_EXE, os.path.join(ROOT, "tests", "continuation_oracle.cpp"), os.path.join(ROOT, "tests", "continuation_cases.H"),
              os.path.join(ROOT, "tests", "compat"), os.path.join(ROOT, "iemic_b200", "csrc", "gmres_guard.h"),
              os.path.join(ROOT, "oracle", "thcm_oracle.h"), os.path.join(ROOT, "iemic_b200", "host"), REF_CONT):
        return ORACLE_EXE
    cmd = ["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "tests", "compat"), "-I", REF_CONT, "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "continuation_oracle.cpp"), "-L", os.path.join(ROOT, "oracle"), "-lthcm_oracle",
           "-Wl,-rpath,$ORIGIN/..", "-o", ORACLE_EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("continuation_oracle.cpp does not compile against the reference's Continuation.H:\n" + r.stderr[-4000:])
    return ORACLE_EXE


GMRES_REF_EXE = os.path.join(ROOT, "oracle", "_ref", "gmres_reference")


def build_gmres_reference():
    """tests/gmres_reference.cpp: the reference's unmodified GMRESSolver.H on plain host vectors (pins oracle/gmres_oracle.cpp).
    -> path of the binary, or None when the reference tree is absent"""
    ref = "/root/reference/src/gmressolver"
    if not os.path.isdir(ref):
        return GMRES_REF_EXE if os.path.exists(GMRES_REF_EXE) else None
    os.makedirs(os.path.dirname(GMRES_REF_EXE), exist_ok=True)
    if _fresh(GMRES_REF_EXE, os.path.join(ROOT, "tests", "gmres_reference.cpp"), os.path.join(ROOT, "tests", "compat"), ref):
        return GMRES_REF_EXE
    # the reference's serial PRINT macro takes one argument but is called with two (GMRESMacros.H:18): define it here;
    # no FMA contraction, like the oracle
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DPRINT(m,v)=do{}while(0)", "-I", os.path.join(ROOT, "tests", "compat"),
           "-I", ref, os.path.join(ROOT, "tests", "gmres_reference.cpp"), "-o", GMRES_REF_EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gmres_reference.cpp does not compile against the reference's GMRESSolver.H:\n" + r.stderr[-4000:])
    return GMRES_REF_EXE
