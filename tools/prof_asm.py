"""ncu driver: a few launches of the assembly kernels (stand-alone Jacobian, stand-alone residual, fused) on the 1-degree grid."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import iemic_b200 as ib
from iemic_b200 import grids
n, m, l = 360, 160, 32
cfg = grids.global_config(n, m, l); mask = grids.synthetic_mask(n, m, l)
G = ib.Ocean(cfg, mask)
for k, v in grids.STARTING_PARS["global"].items(): G.setPar(k, v)
x = torch.from_numpy(grids.synthetic_state(n, m, l, mask)).cuda()
F = torch.empty_like(x)
for _ in range(2):
    G.computeJacobian(x); G.computeRHS(x, F); G.assemble(x, F, jacobian=True)
G.sync()
print("ok", G.launch_count())
