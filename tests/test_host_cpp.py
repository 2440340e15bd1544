"""The C++ host mirror of the reference interfaces (iemic_b200/host/*.H): compiles on the CPU box, runs on the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "iemic_b200", "host")
REF_GMRES = "/root/reference/src/gmressolver"
INC = ["-I", os.path.join(ROOT, "include"), "-I", HOST]


def test_host_mirror_compiles():
    subprocess.run(["g++", "-std=c++17", "-Wall", "-fsyntax-only"] + INC + [os.path.join(HOST, "selftest.cpp")], check=True)
    # the reference builds with -std=c++1y / c++14 (src/CMakeLists.txt:24,30): the mirror, its XML front end included, must too
    subprocess.run(["g++", "-std=c++14", "-Wall", "-fsyntax-only"] + INC + [os.path.join(HOST, "selftest.cpp")], check=True)


@pytest.mark.skipif(not os.path.isdir(REF_GMRES), reason="reference tree not present (GPU box)")
def test_reference_gmressolver_instantiates_on_the_mirror():
    """Drop-in claim, compiled: the reference's UNMODIFIED GMRESSolver<Model,VectorPointer> (src/gmressolver/GMRESSolver.H)
    is instantiated with Model = OceanB200, Vector = DeviceVector.  (Read in place from the reference tree, never copied.)"""
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-DWITH_REFERENCE_GMRES", "-DPRINT(m,v)=do{}while(0)", "-I", os.path.join(ROOT, "tests", "compat"),
           "-I", REF_GMRES] + INC + [os.path.join(HOST, "selftest.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/continuation"), reason="reference tree not present (GPU box)")
def test_reference_continuation_instantiates_on_the_mirror():
    """Drop-in claim for the continuation driver, compiled and linked: the reference's UNMODIFIED Continuation<Model>
    (src/continuation/Continuation.H, 1380 lines: Euler predictor, bordered Newton corrector, computeDFDPar, secant
    destination detection, step control) with Model = std::shared_ptr<OceanB200>.  Teuchos / Utils / JDQZ come from the
    test-only stand-ins in tests/compat."""
    import hostcpp_build
    exe = hostcpp_build.build_continuation_dropin()
    assert exe and os.path.exists(exe)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)   # no GPU here: must fail loudly, not fall back
    assert r.returncode != 0 and "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_reference_continuation_runs_on_gpu():
    """First leg of the reference's integration test src/tests/intt_2dmoc.C:28-40 (3x6x6 2-D MOC box, continuation in
    "Combined Forcing" 0 -> 1 with test/2dmoc/continuation_params.xml), driven by the reference's own Continuation.H on the
    CUDA path; linear solves = restarted GMRES on the device with the device-built point-block Jacobi."""
    import hostcpp_build
    exe = hostcpp_build.build_continuation_dropin()
    if not exe or not os.path.exists(exe):
        pytest.skip("tests/_build/continuation_dropin was not built (needs the reference tree at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "continuation status 0" in r.stdout and "continuation drop-in ok" in r.stdout


@pytest.mark.gpu
def test_reference_intt_2dmoc_chain_on_gpu(tmp_path):
    """All seven legs of src/tests/intt_2dmoc.C:28-120 by the reference's own Continuation.H on the CUDA path, then the
    reference's assertions on the overturning streamfunction of the two end states (:96-97, :127-128)."""
    import numpy as np
    import hostcpp_build
    import moc2d
    from iemic_b200 import grids
    exe = hostcpp_build.build_continuation_dropin()
    if not exe or not os.path.exists(exe):
        pytest.skip("tests/_build/continuation_dropin was not built (needs the reference tree at build time)")
    r = subprocess.run([exe, "chain", str(tmp_path)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "continuation drop-in ok" in r.stdout
    cfg = grids.moc2d_config(moc2d.N_, moc2d.M_, moc2d.L_)
    pmin1, pmax1 = moc2d.psim(cfg, np.fromfile(tmp_path / "state1.bin"))
    pmin2, pmax2 = moc2d.psim(cfg, np.fromfile(tmp_path / "state2.bin"))
    assert abs(pmax1 - 14.7) <= 1e-1 and abs(pmin1) <= 1e-4
    assert abs(pmin1 + pmax2) <= 1e-4 and abs(pmax1 + pmin2) <= 1e-4


@pytest.mark.gpu
def test_reference_test_ocean_continuation_on_gpu(tmp_path):
    """TEST(Ocean, Continuation), src/tests/test_ocean.C:178-223, by the reference's own Continuation.H on the CUDA path
    (8x8x4 North Atlantic box with land, Newton-chord hybrid corrector, five steps; the reference asserts status 0), and the
    same run on the CPU oracle: same point on the branch."""
    import hostcpp_build
    from test_continuation_oracle import _natl8_mask_file, parse_test_ocean
    exe = hostcpp_build.build_continuation_dropin()
    ora = hostcpp_build.build_continuation_oracle()
    if not exe or not os.path.exists(exe) or not ora or not os.path.exists(ora):
        pytest.skip("tests/_build programs were not built (they need the reference tree at build time)")
    mask = _natl8_mask_file(tmp_path)
    r = subprocess.run([exe, "ocean", mask], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    st, par, nx, nf = parse_test_ocean(r.stdout)
    assert st == 0
    o = subprocess.run([ora, "ocean", mask, "gmres"], capture_output=True, text=True, timeout=600)
    assert o.returncode == 0, o.stdout[-3000:]
    st_o, par_o, nx_o, nf_o = parse_test_ocean(o.stdout)
    assert st_o == 0
    assert abs(par - par_o) <= 1e-6 * abs(par_o) and abs(nx - nx_o) <= 1e-6 * nx_o


@pytest.mark.gpu
def test_host_mirror_runs_on_gpu(tmp_path):
    exe = str(tmp_path / "selftest")
    libdir = os.path.join(ROOT, "iemic_b200", "lib")
    subprocess.run(["g++", "-std=c++17", "-O1"] + INC + [os.path.join(HOST, "selftest.cpp"), "-L", libdir, "-liemic_b200",
                                                     "-Wl,-rpath," + libdir, "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "selftest ok" in r.stdout
