"""Pins oracle/gmres_oracle.cpp -- the restatement of GMRESSolver.H every GPU GMRES test is compared with -- to the
reference's own code: tests/gmres_reference.cpp compiles the UNMODIFIED src/gmressolver/GMRESSolver.H (in place, from the
reference tree) on plain host vectors and solves the same systems.  Both are built without FMA contraction and sum in the
same order, so return code, iteration count, final recurrence residual and solution must agree to the last bit."""
import struct
import subprocess

import numpy as np
import pytest

import cases
import hostcpp_build
import oracle_lib as ol


def _reference(tmp_path, rp, col, val, b, m, maxit, tol, binv):
    exe = hostcpp_build.build_gmres_reference()
    if exe is None:
        pytest.skip("reference tree absent and no prebuilt oracle/_ref/gmres_reference")
    n = rp.size - 1
    prob, out = tmp_path / "problem.bin", tmp_path / "out.bin"
    with open(prob, "wb") as f:
        f.write(struct.pack("<5i", n, col.size, m, maxit, 0 if binv is None else 1))
        f.write(struct.pack("<d", tol))
        f.write(np.asarray(rp, np.int32).tobytes())
        f.write(np.asarray(col, np.int32).tobytes())
        f.write(np.asarray(val, np.float64).tobytes())
        f.write(np.asarray(b, np.float64).tobytes())
        if binv is not None:
            f.write(np.asarray(binv, np.float64).tobytes())
    r = subprocess.run([exe, str(prob), str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(out, "rb").read()
    rc, it = struct.unpack_from("<2i", raw, 0)
    res, = struct.unpack_from("<d", raw, 8)
    return rc, it, res, np.frombuffer(raw, np.float64, n, 16)


def _block_inverses(rp, col, val, n):
    blocks = np.tile(np.eye(6), (n // 6, 1, 1))
    rows = np.repeat(np.arange(n), np.diff(rp))
    sel = (col // 6) == (rows // 6)
    dense = np.zeros((n // 6, 6, 6))
    dense[rows[sel] // 6, rows[sel] % 6, col[sel] % 6] = val[sel]
    for k in range(n // 6):
        if np.linalg.matrix_rank(dense[k]) == 6:
            blocks[k] = np.linalg.inv(dense[k])
    return blocks.ravel()


@pytest.mark.parametrize("name,m,maxit,tol,prec,expect_rc", [
    ("natl8", 20, 19, 0.0, False, 1),        # one cycle, fixed length
    ("natl8", 10, 45, 0.0, False, 1),        # restarts (explicit residual between the cycles)
    ("natl8_sres1", 30, 200, 1e-6, True, None), # point-block Jacobi (flexible right preconditioning), stops on the tolerance
    ("rand_periodic", 25, 60, 1e-9, True, None),
    ("2dmoc", 400, 800, 1e-8, True, 0),       # singular Jacobian, consistent right-hand side
])
def test_oracle_gmres_equals_the_reference_solver(tmp_path, name, m, maxit, tol, prec, expect_rc):
    cfg, mask, pars = cases.case(name)
    O = ol.Oracle(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
    x = cases.state(cfg, mask, seed=11)
    val, _ = O.jac(x)
    rp, col = O.graph()
    n = rp.size - 1
    b = ol.spmv(rp, col, val, np.random.default_rng(5).uniform(-1, 1, n))
    binv = _block_inverses(rp, col, val, n) if prec else None
    rc_r, it_r, res_r, x_r = _reference(tmp_path, rp, col, val, b, m, maxit, tol, binv)
    rc_o, x_o, hist, it_o = ol.gmres(rp, col, val, b, np.zeros(n), m, maxit, tol, blockinv=binv)
    assert (rc_o, it_o) == (rc_r, it_r)
    assert it_r > 0
    k = [q for q in range(max(0, it_o - 1), it_o + 1) if hist[q] == res_r]
    assert k, (hist[max(0, it_o - 2):it_o + 2], res_r)   # the final recurrence residual, bit for bit
    if expect_rc is not None:
        assert rc_r == expect_rc
    assert np.array_equal(x_o, x_r), np.abs(x_o - x_r).max()
