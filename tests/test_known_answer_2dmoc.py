"""Known answer I6 of the reference (src/tests/intt_2dmoc.C:96-97, 127-128): after the chain of continuations on the
3x6x6 2-D MOC configuration the overturning streamfunction has psi_max = 14.7 +- 0.1 Sv, psi_min = 0 +- 1e-4, and the
two asymmetric branches mirror each other.  This pins the oracle -- and, through the parity tests, the CUDA path -- to a
number the reference's own test suite asserts, not only to invariants."""
import numpy as np
import pytest

import moc2d
import oracle_lib as ol


class OracleBackend:
    def __init__(self, cfg, mask):
        self.o = ol.Oracle(cfg, mask)

    def set_par(self, k, v):
        self.o.set_par(k, v)

    def get_par(self, k):
        return self.o.get_par(k)

    def rhs(self, x):
        return self.o.rhs(x)

    def jac(self, x):
        return self.o.jac(x)

    def graph(self):
        return self.o.graph()


def test_oracle_reproduces_psi_max_of_intt_2dmoc():
    sym, (pmin1, pmax1), x1 = moc2d.run_branch(OracleBackend, -1)
    assert abs(sym[0] + sym[1]) < 1e-8          # thermally driven state is symmetric about the equator
    assert abs(pmax1 - 14.7) < 1e-1             # EXPECT_NEAR(psiMax1, 14.7, 1e-1)
    assert abs(pmin1) < 1e-4                    # EXPECT_NEAR(psiMin1, 0, 1e-4)
    _, (pmin2, pmax2), x2 = moc2d.run_branch(OracleBackend, +1)
    assert abs(pmin1 + pmax2) < 1e-4            # EXPECT_NEAR(psiMin1, -psiMax2, 1e-4)
    assert abs(pmax1 + pmin2) < 1e-4            # EXPECT_NEAR(psiMax1, -psiMin2, 1e-4)


@pytest.mark.gpu
def test_cuda_path_reproduces_psi_max_of_intt_2dmoc():
    import iemic_b200 as ib

    class GpuBackend:
        def __init__(self, cfg, mask):
            self.g = ib.Ocean(cfg, mask)

        def set_par(self, k, v):
            self.g.setPar(k, v)

        def get_par(self, k):
            return self.g.getPar(k)

        def rhs(self, x):
            return self.g.computeRHS_host(x)

        def jac(self, x):
            return self.g.computeJacobian_host(x)

        def graph(self):
            return self.g.graph()

    sym, (pmin1, pmax1), x1 = moc2d.run_branch(GpuBackend, -1)
    assert abs(pmax1 - 14.7) < 1e-1 and abs(pmin1) < 1e-4
    _, (pmin_o, pmax_o), xo = moc2d.run_branch(OracleBackend, -1)
    assert abs(pmax1 - pmax_o) < 1e-8 and np.abs(x1 - xo).max() < 1e-8 * max(1.0, np.abs(xo).max())
