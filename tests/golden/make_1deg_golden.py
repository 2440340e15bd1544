"""Golden data for BASELINE config 4 (synthetic global 1-degree ocean, 360x160x32): the CPU oracle's residual and Jacobian on
the bench's own state and parameters, reduced to what a test can carry --
  * SHA-256 per model level of everything that must be BIT-IDENTICAL (residual rows U,V,W,P; Jacobian entries outside the
    T,S x T,S blocks; diag(B); the graph),
  * a strided sample of the values that carry a tolerance (tracer residual rows, T,S x T,S Jacobian entries: tanh inside).
Run once on a machine with ~40 GB of RAM (about 4 minutes on one core):  python tests/golden/make_1deg_golden.py
Writes tests/golden/global1deg_golden.npz.  The oracle is test infrastructure (oracle/)."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from iemic_b200 import grids  # noqa: E402

N, M, L = 360, 160, 32
F_STRIDE, J_STRIDE = 53, 499


def level_hashes(arr, owner_level, nlev):
    """sha256 of arr[owner_level == k].tobytes() for every level k"""
    out = []
    order = np.argsort(owner_level, kind="stable")
    bounds = np.searchsorted(owner_level[order], np.arange(nlev + 1))
    for k in range(nlev):
        out.append(hashlib.sha256(np.ascontiguousarray(arr[order[bounds[k]:bounds[k + 1]]]).tobytes()).hexdigest())
    return np.array(out)


def reduce_case(F, val, diagB, rp, col, n, m, l):
    nrows = 6 * n * m * l
    rows = np.repeat(np.arange(nrows, dtype=np.int64), np.diff(rp))
    row_lev = (np.arange(nrows) // (6 * n * m)).astype(np.int32)
    ent_lev = row_lev[rows]
    tr_row = (np.arange(nrows) % 6) >= 4
    tr_ent = ((rows % 6) >= 4) & ((col % 6) >= 4)
    d = {}
    d["F_exact_sha"] = level_hashes(F[~tr_row], row_lev[~tr_row], l)
    d["B_sha"] = level_hashes(diagB, row_lev, l)
    d["J_exact_sha"] = level_hashes(val[~tr_ent], ent_lev[~tr_ent], l)
    d["graph_sha"] = np.array([hashlib.sha256(rp.astype(np.int32).tobytes()).hexdigest(), hashlib.sha256(col.astype(np.int32).tobytes()).hexdigest()])
    fi = np.flatnonzero(tr_row)[::F_STRIDE]
    ji = np.flatnonzero(tr_ent)[::J_STRIDE]
    d["F_idx"], d["F_val"] = fi.astype(np.int64), F[fi]
    d["J_idx"], d["J_val"] = ji.astype(np.int64), val[ji]
    rowmax = np.maximum.reduceat(np.abs(val), rp[:-1])
    d["J_rowscale"] = rowmax[rows[ji]]
    d["norms"] = np.array([np.sqrt(np.dot(F, F)), np.abs(F).max(), np.sqrt(np.dot(val, val))])
    return d


if __name__ == "__main__":
    cfg = grids.global_config(N, M, L)
    mask = grids.synthetic_mask(N, M, L)
    O = ol.Oracle(cfg, mask)
    for k, v in grids.STARTING_PARS["global"].items():
        O.set_par(k, v)
    x = grids.synthetic_state(N, M, L, mask)          # the bench's state
    F = O.rhs(x)
    val, diagB = O.jac(x)
    assert O.dropped() == 0
    rp, col = O.graph()
    d = reduce_case(F, val, diagB, rp, col, N, M, L)
    np.savez_compressed(os.path.join(HERE, "global1deg_golden.npz"), **d)
    print("written", {k: (v.shape, v.dtype) for k, v in d.items()})
