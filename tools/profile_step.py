"""Short driver for ncu: a few launches of every hot kernel on the bench workload (1-degree grid by default)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import iemic_b200 as ib  # noqa: E402
from iemic_b200 import grids  # noqa: E402

grid = sys.argv[1] if len(sys.argv) > 1 else "360x160x32"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
n, m, l = [int(v) for v in grid.split("x")]
cfg = grids.global_config(n, m, l)
mask = grids.synthetic_mask(n, m, l)
G = ib.Ocean(cfg, mask)
for k, v in grids.STARTING_PARS["global"].items():
    G.setPar(k, v)
x = torch.from_numpy(grids.synthetic_state(n, m, l, mask)).cuda()
F = torch.empty_like(x)
y = torch.empty_like(x)
for _ in range(reps):
    G.computeJacobian(x)
    G.computeRHS(x, F)
    G.assemble(x, F, jacobian=True)   # the step's own kernel: Jacobian + residual fused
    G.applyMatrix(x, y)
G.sync()
# row N3: scaling and the point-block-Jacobi builder on the current Jacobian
binv = torch.empty(G.nrows * 6, dtype=torch.float64, device="cuda")
rs = torch.empty(G.nrows, dtype=torch.float64, device="cuda")
ri = torch.empty_like(rs)
G.block_jacobi(binv)
G.recompute_scaling(rs, None)
G.reciprocal(rs, ri)
G.left_scale(rs)
G.left_scale(ri)
G.computeJacobian(x)
del binv, rs, ri
b = y.clone()
xs = torch.zeros_like(x)
G.solve(b, xs, m=8, maxit=7, tol=0.0, ortho=ib.CGS2F, check_every=0)
G.solve(b, xs, m=4, maxit=3, tol=0.0, ortho=ib.MGS, check_every=0)
G.sync()
print("ok", G.launch_count())
