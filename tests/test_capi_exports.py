"""The C-ABI library loads and exports every function include/iemic_b200.h declares (no compute calls: no GPU here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "iemic_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:thcm_b200|gmres_b200|vec_b200|iemic_b200|coupling_b200)_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from iemic_b200 import build, capi
    build.build()
    L = capi.lib()
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), n
    assert set(capi.EXPORTED) == set(names)


def test_fortran_abi_library_exports_every_declared_symbol():
    """Seam S1 (include/thcm_fortran_abi.h): the Fortran-named symbols THCM.C binds (src/ocean/THCM.C:47-174)."""
    import ctypes as C
    from iemic_b200 import build
    import fortran_abi
    build.build()
    L = C.CDLL(fortran_abi.LIB)
    names = fortran_abi.declared()
    assert len(names) >= 15 and "rhs_" in names and "matrix_" in names and "__m_mat_MOD_set_pointers" in names
    for n in names:
        assert hasattr(L, n), n


def test_no_gpu_fails_loudly():
    """Without a CUDA device creation must fail with an error code and message, never fall back to the CPU."""
    import ctypes as C
    import numpy as np
    from iemic_b200 import capi, grids
    L = capi.lib()
    if L.iemic_b200_device_count() > 0:
        pytest.skip("a GPU is present")
    cfg = grids.default_config(8, 8, 4)
    with pytest.raises(capi.B200Error) as e:
        capi.Ocean(cfg, grids.no_land_mask(8, 8, 4, False))
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under iemic_b200/ or include/ may reference it."""
    bad = []
    for base in ("iemic_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if os.sep + "build" in dp or os.sep + "lib" in dp:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".H")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if "oracle_lib" in txt or "thcm_oracle" in txt or "libthcm_oracle" in txt:
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
