"""The reference's XML parameter files, read unchanged (iemic_b200/params.py), against what the test-suite hard-codes: the
Python cases (tests/cases.py, iemic_b200/grids.py) and the C++ set-ups of the Continuation harness
(tests/continuation_cases.H, printed by `continuation_oracle cases`).  The reference tree only exists in the build container."""
import math
import os
import subprocess

import pytest

import cases
from iemic_b200 import params

REF = "/root/reference"
needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "test", "ocean")), reason="reference checkout not present")


def _same_cfg(a, b):
    assert set(a) == set(b), set(a) ^ set(b)
    for k in a:
        assert a[k] == b[k], (k, a[k], b[k])


@needs_ref
@pytest.mark.parametrize("xml,case,mask", [
    ("test/ocean/ocean_params.xml", "natl8", "mask_natl8"),
    ("test/2dmoc/ocean_params.xml", "2dmoc", None),
    ("test/matrix/ocean_params.xml", None, "test6x6x4"),
])
def test_thcm_config_from_the_reference_files(xml, case, mask):
    cfg, start, mask_file = params.thcm_config(params.read_parameter_list(os.path.join(REF, xml)))
    assert mask_file == mask
    if case:
        c2, _, _ = cases.case(case)
        _same_cfg(cfg, c2)
    assert start["SPL2"] == 0.01 and "Combined Forcing" in start


@needs_ref
def test_reft_ocean_config_is_the_golden_state_case():
    cfg, start, mask_file = params.thcm_config(params.read_parameter_list(os.path.join(REF, "test/ocean/reft_ocean_params.xml")))
    _same_cfg(cfg, cases.reference_state()[0])   # "Mixing" = 2 maps to the always-on operator, see params.thcm_config
    assert mask_file == "mask_gateway" and start["Salinity Forcing"] == 0.1


@needs_ref
def test_unsupported_options_fail_loudly():
    with pytest.raises(ValueError, match="idealized forcing"):      # Levitus data files
        params.thcm_config(params.read_parameter_list(os.path.join(REF, "run/ocean/global/ocean_params.xml")))
    with pytest.raises(ValueError, match="not implemented"):         # salinity perturbation mask
        params.thcm_config(params.read_parameter_list(os.path.join(REF, "run/2dmoc/ocean_params.xml")))
    with pytest.raises(ValueError, match="unknown THCM parameters"):  # stale names: the reference's validation rejects them too
        params.thcm_config(params.read_parameter_list(os.path.join(REF, "parameterfiles/ocean_params.xml")))
    with pytest.raises(ValueError, match="type"):
        params.thcm_config({"THCM": {"Global Grid-Size n": "16"}})


def test_defaults_without_a_file():
    cfg, start, mask_file = params.thcm_config({})
    assert (cfg["n"], cfg["m"], cfg["l"]) == (16, 16, 16) and cfg["rho_mixing"] == 1 and not cfg["periodic"]
    assert cfg["xmin"] == 286.0 * math.pi / 180.0 and start == {} and mask_file is None


def _harness_cases():
    import hostcpp_build
    exe = hostcpp_build.build_continuation_oracle()
    if exe is None:
        pytest.skip("oracle/_ref/continuation_oracle not built")
    out = subprocess.run([exe, "cases"], capture_output=True, text=True, check=True).stdout
    res, cur = {}, None
    for line in out.splitlines():
        if line.startswith("CASE "):
            cur = res.setdefault(line[5:], {"cfg": {}, "cont": {}})
        elif line.startswith("cfg "):
            _, k, v = line.split(" ", 2)
            cur["cfg"][k] = float(v)
        elif line.startswith("cont "):
            k, v = line[5:].split(" = ", 1)
            cur["cont"][k] = v
    return res


@needs_ref
@pytest.mark.parametrize("name,ocean_xml,cont_xml", [
    ("test_ocean", "test/ocean/ocean_params.xml", "test/ocean/continuation_params.xml"),
    ("reft_ocean", "test/ocean/reft_ocean_params.xml", "test/ocean/reft_continuation_params.xml"),
    ("2dmoc", "test/2dmoc/ocean_params.xml", "test/2dmoc/continuation_params.xml"),
])
def test_harness_cases_are_the_reference_files(name, ocean_xml, cont_xml):
    h = _harness_cases()[name]
    cfg, _, _ = params.thcm_config(params.read_parameter_list(os.path.join(REF, ocean_xml)))
    assert set(h["cfg"]) == set(cfg)
    for k, v in cfg.items():
        assert h["cfg"][k] == float(v), (k, h["cfg"][k], v)   # bit for bit, bounds included
    cont = params.read_parameter_list(os.path.join(REF, cont_xml))
    assert set(h["cont"]) == set(cont), set(h["cont"]) ^ set(cont)
    for k, v in cont.items():
        got = h["cont"][k]
        if isinstance(v, bool):
            assert got == ("1" if v else "0"), k
        elif isinstance(v, (int, float)):
            assert float(got) == float(v), (k, got, v)
        else:
            assert got == v, (k, got, v)


# ---- the same front end in C++ (iemic_b200/host/ParameterListXML.H), through tests/xml_frontend.cpp -------------------------
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def xml_frontend(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("xml") / "xml_frontend")
    subprocess.run(["g++", "-std=c++17", "-Wall", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "iemic_b200", "host"),
                    os.path.join(ROOT, "tests", "xml_frontend.cpp"), "-o", exe], check=True)
    return exe


@needs_ref
@pytest.mark.parametrize("xml,shape", [("test/ocean/ocean_params.xml", (8, 8, 4)), ("test/2dmoc/ocean_params.xml", (3, 6, 6)),
                                       ("test/ocean/reft_ocean_params.xml", (16, 16, 16)), ("test/matrix/ocean_params.xml", (6, 6, 4))])
def test_cpp_front_end_equals_the_python_one(xml_frontend, tmp_path, xml, shape):
    import numpy as np
    from iemic_b200 import checkpoint
    path = os.path.join(REF, xml)
    cfg, start, mask_file = params.thcm_config(params.read_parameter_list(path))
    out_mask = str(tmp_path / "mask.bin")
    r = subprocess.run([xml_frontend, path, os.path.join(REF, "data", "mkmask"), out_mask], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    got_cfg, got_start, got_mask = {}, {}, None
    for line in r.stdout.splitlines():
        if line.startswith("cfg "):
            _, k, v = line.split(" ", 2)
            got_cfg[k] = float(v)
        elif line.startswith("start "):
            k, v = line[6:].split(" = ")
            got_start[k] = float(v)
        elif line.startswith("mask "):
            got_mask = line[5:] or None
    assert (cfg["n"], cfg["m"], cfg["l"]) == shape
    assert got_cfg == {k: float(v) for k, v in cfg.items()}
    assert got_start == start and got_mask == mask_file
    if mask_file:
        n, m, l = shape
        a = checkpoint.read_mask(os.path.join(REF, "data", "mkmask", mask_file), n, m, l)
        assert np.array_equal(np.fromfile(out_mask, np.int32).reshape(a.shape), a)


@needs_ref
def test_cpp_front_end_rejects_what_the_python_one_rejects(xml_frontend):
    for xml, msg in (("run/ocean/global/ocean_params.xml", "idealized forcing"), ("run/2dmoc/ocean_params.xml", "not implemented"),
                     ("parameterfiles/ocean_params.xml", "unknown THCM parameter")):
        r = subprocess.run([xml_frontend, os.path.join(REF, xml)], capture_output=True, text=True)
        assert r.returncode == 3 and msg in r.stdout, (xml, r.stdout)


@needs_ref
def test_cpp_reader_sees_every_parameter(xml_frontend):
    """all parameter files under test/ and run/ of the reference: same names, types and values as the Python reader"""
    import glob

    def flat(d, prefix=""):
        out = {}
        for k, v in d.items():
            if isinstance(v, dict):
                out.update(flat(v, prefix + k + "/"))
            else:
                out[prefix + k] = v
        return out

    files = sorted(glob.glob(os.path.join(REF, "test", "*", "*.xml")) + glob.glob(os.path.join(REF, "run", "*", "*.xml")))
    assert len(files) > 40
    checked = 0
    for f in files:
        try:
            want = flat(params.read_parameter_list(f))
        except Exception:
            continue   # not a ParameterList file (or a type neither reader handles)
        r = subprocess.run([xml_frontend, "--list", f], capture_output=True, text=True)
        assert r.returncode == 0, (f, r.stdout)
        got = {}
        for line in r.stdout.splitlines():
            k, rest = line.split(" = ", 1)
            typ, _, val = rest.partition(" ")
            got[k] = {"bool": lambda s: bool(int(s)), "int": int, "double": float, "char": str, "string": str}[typ](val)
        assert set(got) == set(want), (f, set(got) ^ set(want))
        for k in want:
            assert got[k] == want[k] and type(got[k]) is type(want[k]), (f, k, got[k], want[k])
        checked += 1
    assert checked > 40


@needs_ref
def test_ocean_setup_reads_files_and_masks():
    import numpy as np
    from iemic_b200 import grids
    from iemic_b200.capi import PAR_NAMES
    cfg, mask, start = params.ocean_setup(os.path.join(REF, "test/ocean/ocean_params.xml"), os.path.join(REF, "data", "mkmask"))
    c2, m2, _ = cases.case("natl8")
    _same_cfg(cfg, c2)
    assert np.array_equal(mask, np.asarray(m2).reshape(mask.shape))
    assert all(k in PAR_NAMES for k in start)      # Ocean.from_xml passes these names to setPar
    cfg, mask, start = params.ocean_setup(os.path.join(REF, "test/2dmoc/ocean_params.xml"))   # Topography 1, no mask file
    assert np.array_equal(mask, grids.no_land_mask(3, 6, 6, True))
    assert start["Horizontal Ekman-Number"] == 371.764

