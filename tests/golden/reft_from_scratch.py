"""The reference's golden-vector regression (src/tests/reft_ocean.C:31-80) from the ZERO state, on the CPU oracle:
natural-parameter continuation in Combined Forcing, 0 -> the value stored in test/ocean/ocean_reference.h5 (0.0200002), Newton
with a sparse LU at every step, then the comparison of the end state with the golden vector.  Six minutes on one core, so it
is a script and not a test; tests/test_reference_state.py runs the short form (Newton from the golden state, same root).

Output of the run recorded in tests/test_reference_state.py and DESIGN.md:
    u  norm 0.0979067 vs 0.0979069 (reference bound 1e-3)   max |diff| 1.2e-7
    v  norm 0.0224026 vs 0.0224030                          max |diff| 9.8e-8
    T  norm 3.5162672 vs 3.5162058                          max |diff| 4.1e-6
    S  norm 0.0351627 vs 0.0351621                          max |diff| 4.2e-8
    w                                                       max |diff| 1.1e-6
    p  (defined up to its gauge)                            max |diff| 1.7e-3

usage: python tests/golden/reft_from_scratch.py"""
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), os.path.dirname(HERE)]
import cases          # noqa: E402
import oracle_lib as ol  # noqa: E402

COMB = 19


def main():
    cfg, mask, pars, xg = cases.reference_state()
    O = ol.Oracle(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
    rp, col = O.graph()
    n = rp.size - 1
    shift = np.zeros(n)
    shift[3::6] = 1e-9   # pressure gauge: regularises the constant-pressure null vector
    print("unknowns", n, " ||F(golden state)|| =", np.linalg.norm(O.rhs(xg)))

    def newton_update(x, F):
        val, _ = O.jac(x)
        J = sp.csr_matrix((val, col, rp), shape=(n, n)).tocsc() + sp.diags(shift).tocsc()
        return spl.splu(J).solve(-F)

    target = pars[COMB]
    x, v, step, t0 = np.zeros(n), 0.0, target / 8, time.time()
    while v < target:
        vn = min(v + step, target)
        O.set_par(COMB, vn)
        xn, ok = x.copy(), False
        for it in range(12):
            F = O.rhs(xn)
            nf = np.linalg.norm(F)
            if nf < 1e-9:
                ok = True
                break
            xn = xn + newton_update(xn, F)
        print("Combined Forcing %.7f: %d Newton iterations, ||F|| = %.2e, %s, %.0f s" % (vn, it, nf, "ok" if ok else "FAILED", time.time() - t0), flush=True)
        if not ok:
            step /= 2
            O.set_par(COMB, v)
            continue
        x, v = xn, vn
    X, G = x.reshape(-1, 6), xg.reshape(-1, 6)
    for f, name in enumerate("uvwpTS"):
        print("%s  norm %.7f vs %.7f   max |diff| %.2e" % (name, np.linalg.norm(X[:, f]), np.linalg.norm(G[:, f]), np.abs(X[:, f] - G[:, f]).max()))


if __name__ == "__main__":
    main()
