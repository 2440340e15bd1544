"""Seam S1 link test: every Fortran symbol the reference's C++ layer binds (extern "C" block of src/ocean/THCM.C:47-174, GNU name
mangling of src/utils/my_f2c.H:15-42) must be exported by libthcm_b200_fortran.so, so that an unmodified THCM.C links against it
instead of the Fortran objects.  Two forms: (1) with the reference tree at hand, the extern block itself is compiled into a
translation unit that takes the address of every declared routine and is linked with -Wl,--no-undefined; (2) everywhere, the
committed list of mangled names (tests/golden/thcm_fortran_symbols.txt, made by this file's __main__) is checked with nm."""
import os
import re
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference/src/ocean/THCM.C"
LIB = os.path.join(ROOT, "iemic_b200", "lib", "libthcm_b200_fortran.so")
LIST = os.path.join(HERE, "golden", "thcm_fortran_symbols.txt")


def extern_block():
    src = open(REF).read()
    a = src.index('extern "C" {')
    b = src.index("}//extern", a)
    return src[a:b + 1]


def mangled_names(block):
    names = ["%s_" % n for n in re.findall(r"_SUBROUTINE_\((\w+)\)", re.sub(r"_MODULE_SUBROUTINE_\([^)]*\)", "", block))]
    names += ["__%s_MOD_%s" % (m, f) for m, f in re.findall(r"_MODULE_SUBROUTINE_\(\s*(\w+)\s*,\s*(\w+)\s*\)", block)]
    return sorted(set(names))


def exported():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if " T " in ln}


def test_every_bound_symbol_is_exported():
    want = [ln.strip() for ln in open(LIST) if ln.strip()]
    assert len(want) >= 60
    missing = sorted(set(want) - exported())
    assert not missing, missing


@pytest.mark.skipif(not os.path.exists(REF), reason="needs the reference tree (build container only)")
def test_extern_block_links_without_undefined_symbols(tmp_path):
    block = extern_block()
    names = mangled_names(block)
    assert names == [ln.strip() for ln in open(LIST) if ln.strip()], "tests/golden/thcm_fortran_symbols.txt is stale: python tests/test_fortran_link.py"
    tu = tmp_path / "thcm_extern.cpp"
    tu.write_text("#define GNU\n#include \"%s\"\n%s\nvoid* thcm_symbol_table[] = {%s};\nint main() { return thcm_symbol_table[0] == 0; }\n"
                  % ("/root/reference/src/utils/my_f2c.H", block, ", ".join("(void*)&%s" % n for n in names)))
    exe = tmp_path / "thcm_extern"
    r = subprocess.run(["g++", "-std=c++14", str(tu), "-o", str(exe), "-Wl,--no-undefined", "-L" + os.path.dirname(LIB), "-lthcm_b200_fortran",
                        "-liemic_b200", "-Wl,-rpath," + os.path.dirname(LIB)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


if __name__ == "__main__":
    with open(LIST, "w") as f:
        f.write("\n".join(mangled_names(extern_block())) + "\n")
    print("written", LIST)
