"""The reference's own integration test of this path, src/tests/intt_2dmoc.C:28-137, run by the reference's UNMODIFIED
Continuation<Model> (src/continuation/Continuation.H, compiled in place) on the CPU oracle: seven pseudo-arclength legs on the
3x6x6 2-D MOC box, then the assertions the reference makes on the overturning streamfunction of the two end states
(psi_max = 14.7 +- 0.1 Sv, psi_min = 0 +- 1e-4, the second state the mirror image of the first within 1e-4).  This pins the
oracle's F and J (invariant I6) through the reference's own driver rather than a solver of ours.

Two linear solvers: a dense LU (exact up to the pressure null vector), and a host walk-through of the product's GMRES driver
(CGS2 + the accuracy guard of iemic_b200/csrc/gmres_guard.h).  Without the guard the CGS2 driver fails the first or the second leg: the
Jacobian is singular here (no fixed pressure points), Newton's last right-hand sides sit at the rounding floor of F, and a
solve accepted on the recurrence residual returns a multiple of the null vector (same failure seen on the GPU)."""
import os
import subprocess

import numpy as np
import pytest

import hostcpp_build
import moc2d
from iemic_b200 import grids


def _exe():
    exe = hostcpp_build.build_continuation_oracle()
    if exe is None:
        pytest.skip("reference tree absent and no prebuilt oracle/_ref/continuation_oracle")
    return exe


def _run(tmp_path, mode, nlegs, env=None, timeout=600):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([_exe(), str(tmp_path), mode, str(nlegs)], capture_output=True, text=True, env=e, timeout=timeout)
    return r.returncode, r.stdout


def test_intt_2dmoc_chain_with_the_reference_continuation(tmp_path):
    rc, out = _run(tmp_path, "lu", 7)
    assert rc == 0 and "continuation on the oracle ok" in out, out[-2000:]
    cfg = grids.moc2d_config(moc2d.N_, moc2d.M_, moc2d.L_)
    x1 = np.fromfile(tmp_path / "state1.bin")
    x2 = np.fromfile(tmp_path / "state2.bin")
    pmin1, pmax1 = moc2d.psim(cfg, x1)
    pmin2, pmax2 = moc2d.psim(cfg, x2)
    assert abs(pmax1 - 14.7) <= 1e-1   # intt_2dmoc.C:96
    assert abs(pmin1) <= 1e-4          # :97
    assert abs(pmin1 + pmax2) <= 1e-4  # :127
    assert abs(pmax1 + pmin2) <= 1e-4  # :128


def test_product_gmres_driver_needs_the_accuracy_guard(tmp_path):
    import re
    # the chain's first two legs with the walk-through of the product's GMRES driver (guard on)
    rc, out = _run(tmp_path, "cgs2", 2)
    assert rc == 0, out[-2000:]
    legs = [ln for ln in out.splitlines() if ln.startswith("leg ")]
    assert len(legs) == 2 and all("status 0" in ln for ln in legs), out[-2000:]
    # one solve at the rounding floor of F at the end of leg 1, guard on and off.  Off: the recurrence residual "converges"
    # (rc 0) to an update that is a multiple of the null vector, nine orders of magnitude larger than the guarded one.  (With
    # ORACLE_GUARD=0 the chain itself aborts at the minimum step size in leg 1 or 2, depending on the last bits of the set-up.)
    rc, out = _run(tmp_path, "guardcheck", 1)
    assert rc == 0, out[-2000:]
    res = {int(g): (int(r), float(nb), float(nx)) for g, r, nb, nx in
           re.findall(r"guardcheck: guard (\d), rc (\d+), \|\|b\|\| = (\S+), \|\|x\|\| = (\S+)", out)}
    assert set(res) == {0, 1}, out[-2000:]
    (rc_on, nb, x_on), (rc_off, _, x_off) = res[1], res[0]
    assert nb < 1e-10                      # F at its rounding floor
    # b is inconsistent beyond the tolerance, so no solver can honestly return 0 here.  Guard on: "accuracy limit" (2; 1 if the
    # iteration merely stagnated) with an update of the order of ||b||.  Guard off: never a small update reported as converged
    # (on this machine: rc 0 with ||x|| = 0.19).
    assert rc_on in (1, 2) and x_on <= 1e4 * nb
    assert not (rc_off == 0 and x_off <= 1e4 * nb)

def _natl8_mask_file(tmp_path):
    import cases
    cfg, mask, _ = cases.case("natl8")
    path = tmp_path / "mask_natl8.bin"
    np.asarray(mask, np.int32).tofile(path)
    return str(path)


def parse_test_ocean(out):
    """-> (status, parameter value, ||x||, ||F||) of the line continuation_cases.H prints"""
    import re
    m = re.search(r"test_ocean continuation: status (-?\d+), Combined Forcing = (\S+), \|\|x\|\| = (\S+), \|\|F\|\| = (\S+)", out)
    assert m, out[-2000:]
    return int(m.group(1)), float(m.group(2)), float(m.group(3)), float(m.group(4))


def test_test_ocean_continuation_with_the_reference_continuation(tmp_path):
    """TEST(Ocean, Continuation), src/tests/test_ocean.C:178-223 (8x8x4 North Atlantic box, Newton-chord hybrid corrector,
    five steps): the reference asserts status 0.  Run with the oracle's GMRESSolver restatement and with the walk-through of
    the product's GMRES driver: same branch point."""
    mask = _natl8_mask_file(tmp_path)
    res = {}
    for mode in ("gmres", "cgs2"):
        r = subprocess.run([_exe(), "ocean", mask, mode], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:]
        res[mode] = parse_test_ocean(r.stdout)
        assert res[mode][0] == 0
        assert res[mode][1] > 0.0 and res[mode][2] > 0.0
    assert abs(res["gmres"][1] - res["cgs2"][1]) <= 1e-7 * abs(res["gmres"][1])
    assert abs(res["gmres"][2] - res["cgs2"][2]) <= 1e-7 * res["gmres"][2]


REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "test", "2dmoc")), reason="reference checkout not present")
def test_run_ocean_flow_from_the_reference_files(tmp_path):
    """The flow of the reference's executable src/main/run_ocean.C:44-75 (without the eigen-solver): ocean_params.xml,
    continuation_params.xml and the land-mask file read unchanged (iemic_b200/host/ParameterListXML.H), then the reference's
    Continuation::run().  Same results as the hard-coded set-ups of the other tests, to the last printed digit."""
    import re
    exe = _exe()
    masks = os.path.join(REF, "data", "mkmask")
    r = subprocess.run([exe, "xml", os.path.join(REF, "test/2dmoc/ocean_params.xml"), os.path.join(REF, "test/2dmoc/continuation_params.xml"),
                        masks, "lu"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    m = re.search(r"run_ocean: status 0, Combined Forcing = (\S+), \|\|x\|\| = (\S+),", r.stdout)
    assert m, r.stdout[-2000:]
    rc, out = _run(tmp_path, "lu", 1)
    h = re.search(r"leg Combined Forcing -> 1: status 0, value (\S+), \|\|x\|\| = (\S+),", out)
    assert rc == 0 and h, out[-2000:]
    assert m.group(1) == h.group(1) and abs(float(m.group(2)) - float(h.group(2))) <= 1e-6 * float(h.group(2))
    # test/ocean: 8x8x4 box with the land mask file of data/mkmask
    r = subprocess.run([exe, "xml", os.path.join(REF, "test/ocean/ocean_params.xml"), os.path.join(REF, "test/ocean/continuation_params.xml"),
                        masks, "gmres"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    m = re.search(r"run_ocean: status 0, Combined Forcing = (\S+), \|\|x\|\| = (\S+),", r.stdout)
    o = subprocess.run([exe, "ocean", _natl8_mask_file(tmp_path), "gmres"], capture_output=True, text=True, timeout=600)
    st, par, nx, nf = parse_test_ocean(o.stdout)
    assert st == 0 and float(m.group(1)) == par and float(m.group(2)) == nx
