"""Times jac_kernel / rhs_kernel of the library selected by IEMIC_B200_LIB (kernel-variant experiments)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import iemic_b200 as ib
from iemic_b200 import grids
n, m, l = [int(v) for v in os.environ.get('GRID', '360x160x32').split('x')]
vm = int(os.environ.get("VMIX", "1")); kind = os.environ.get("MASK", "global")
cfg = grids.global_config(n, m, l, vmix=vm); mask = grids.synthetic_mask(n, m, l, kind=kind)
G = ib.Ocean(cfg, mask)
for k, v in grids.STARTING_PARS["global"].items(): G.setPar(k, v)
x = torch.from_numpy(grids.synthetic_state(n, m, l, mask, kind=os.environ.get("STATE", "random"))).cuda()
F = torch.empty_like(x)
for _ in range(3): G.computeJacobian(x); G.computeRHS(x, F); G.assemble(x, F, jacobian=True)
torch.cuda.synchronize()
def t(fn, reps=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
y = torch.empty_like(x)
sp = t(lambda: G.applyMatrix(x, y))
print(os.path.basename(os.environ.get("IEMIC_B200_LIB", "default")), "spmv ms %.4f" % sp, "vmix", vm, "mask", kind, "state", os.environ.get("STATE", "random"),
      "jac ms %.4f rhs ms %.4f fused ms %.4f" % (t(lambda: G.computeJacobian(x)), t(lambda: G.computeRHS(x, F)), t(lambda: G.assemble(x, F, jacobian=True))))
