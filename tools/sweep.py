"""BASELINE config 5: CSR SpMV + GMRES(m) orthogonalisation sweep over grid sizes against the HBM roofline (one GPU).

For every grid: Jacobian assembled on the device from a synthetic state, then SpMV alone and one GMRES(m) cycle
(fixed iteration count, identity preconditioner) per orthogonalisation mode, timed with CUDA events.
Prints one JSON line per grid; `--out FILE` also writes them to FILE.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import iemic_b200 as ib  # noqa: E402
from iemic_b200 import grids  # noqa: E402
import bench  # noqa: E402  (alg_bytes, peaks)

GRIDS = ["96x38x12", "180x80x16", "360x160x16", "360x160x32", "720x320x16", "720x320x32"]


def timed(fn, reps, warm):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grids", default=",".join(GRIDS))
    ap.add_argument("--m", type=int, default=30)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    peak, kind = bench.peaks()
    lines = []
    for g in args.grids.split(","):
        n, m, l = [int(v) for v in g.split("x")]
        cfg = grids.global_config(n, m, l)
        mask = grids.synthetic_mask(n, m, l, kind="global")
        G = ib.Ocean(cfg, mask)
        for k, v in grids.STARTING_PARS["global"].items():
            G.setPar(k, v)
        x = grids.synthetic_state(n, m, l, mask)
        gid = G.global_ids()
        dx = torch.from_numpy(np.where(gid >= 0, x[np.maximum(gid, 0)], 0.0)).cuda()
        dF = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        ms_asm = timed(lambda: G.assemble(dx, dF, jacobian=True), 10, 3)
        ncell, nrows, nnz = n * m * l, 6 * n * m * l, G.nnz
        jac_b, rhs_b, spmv_b, cycle_b = bench.alg_bytes(ncell, nrows, nnz, args.m)
        ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
        y = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        ms_spmv = timed(lambda: G.applyMatrix(ones, y), 30, 5)
        b = y.clone()
        xs = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        out = {"grid": g, "unknowns": nrows, "nnz": nnz, "assembly_ms": ms_asm, "assembly_frac": (jac_b + rhs_b) / (ms_asm * 1e-3) / 1e9 / peak,
               "spmv_ms": ms_spmv, "spmv_gbs": spmv_b / (ms_spmv * 1e-3) / 1e9, "spmv_frac": spmv_b / (ms_spmv * 1e-3) / 1e9 / peak,
               "peak_gbs": peak, "peak_kind": kind, "m": args.m}
        for name, mode in (("mgs", ib.MGS), ("cgs2", ib.CGS2), ("cgs2f", ib.CGS2F)):
            def cycle():
                xs.zero_()
                G.solve(b, xs, m=args.m, maxit=args.m - 1, tol=0.0, ortho=mode, check_every=0)
            ms = timed(cycle, 3, 1)
            out["gmres_%s_iters_per_s" % name] = args.m / (ms * 1e-3)
            out["gmres_%s_frac" % name] = cycle_b / (ms * 1e-3) / 1e9 / peak  # CGS2 byte model for all modes
        G.close()
        del dx, dF, ones, y, b, xs
        torch.cuda.empty_cache()
        print(json.dumps(out), flush=True)
        lines.append(out)
    if args.out:
        with open(args.out, "w") as f:
            for o in lines:
                f.write(json.dumps(o) + "\n")


if __name__ == "__main__":
    main()
