"""Times the SpMV variants (IEMIC_SPMV_VARIANT) on the bench workload; run once per variant in a fresh process."""
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
G.computeJacobian(x)
y = torch.empty_like(x)
for _ in range(5): G.applyMatrix(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): G.applyMatrix(x, y)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
b = 12 * G.nnz + 20 * G.nrows
print("variant", os.environ.get("IEMIC_SPMV_VARIANT"), "rvar", os.environ.get("IEMIC_SPMV_RVAR"), "regular", os.environ.get("IEMIC_SPMV_REGULAR"), "ms %.4f GB/s %.0f" % (ms, b / ms / 1e6), "checksum %.17g" % float(y.sum()))
