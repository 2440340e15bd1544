"""ncu driver: a few launches of the SpMV on the 1-degree grid."""
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
y = torch.empty_like(x)
G.computeJacobian(x)
for _ in range(4): G.applyMatrix(x, y)
G.sync()
print("ok", G.launch_count())
