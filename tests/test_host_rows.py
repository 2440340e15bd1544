"""The product's __host__ __device__ row code (the arithmetic the CUDA kernels run), compiled for the CPU by the test
harness tests/host_rowcheck.cpp, against the structure-faithful oracle: bit-exact on every case."""
import numpy as np
import pytest

import cases
import hostcheck_lib as hl
import oracle_lib as ol

CASES = ["natl8", "natl8_sres1", "test6x6x4", "2dmoc", "rand_periodic", "rand_box", "rand_fixp", "global_half", "coupled4"]


def _pair(name):
    cfg, mask, pars = cases.case(name)
    O, H = ol.Oracle(cfg, mask), hl.HostCheck(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
        H.set_par(k, v)
    return cfg, mask, O, H


@pytest.mark.parametrize("name", CASES)
def test_rows_bit_exact(name):
    cfg, mask, O, H = _pair(name)
    rp, col = O.graph()
    rp2, col2 = H.graph()
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2)
    O.rhs(np.zeros(O.ndim))   # the oracle masks Frc inside `boundaries`, i.e. during an evaluation
    assert np.array_equal(O.frc(), H.frc())
    for kind, seed in (("random", 1), ("smooth", 2)):
        x = cases.state(cfg, mask, seed=seed, kind=kind)
        assert np.array_equal(O.rhs(x), H.rhs(x))
        vo, bo = O.jac(x)
        vh, bh = H.jac(x)
        assert O.dropped() == 0
        assert np.array_equal(vo, vh)
        assert np.array_equal(bo, bh)


def test_unsupported_configurations_fail_loudly():
    cfg, mask, pars = cases.case("natl8")
    with pytest.raises(RuntimeError):
        hl.HostCheck(dict(cfg, vmix=2), mask)
    with pytest.raises(RuntimeError):
        hl.HostCheck(dict(cfg, coupled_S=1), mask)
    H = hl.HostCheck(cfg, mask)
    with pytest.raises(RuntimeError):
        H.set_par("MKAP", 1.0)


def test_decomp2d_rule():
    from iemic_b200 import grids
    assert grids.decomp2d(360, 160, 8) == (4, 2)
    assert grids.decomp2d(360, 160, 4) == (4, 1)   # r <= r_min keeps scanning: |160/1 - 360/4| = 70 beats 2x2 (100)
    assert grids.decomp2d(360, 160, 2) == (2, 1)
    assert grids.decomp2d(96, 38, 8) == (4, 2)
    assert grids.decomp2d(96, 38, 1) == (1, 1)
