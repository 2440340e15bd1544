#!/usr/bin/env python
"""bench.py -- throughput of the B200 hot path on BASELINE.json's workload.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU algorithm (oracle port) on host cores

Workload (config.workload): BASELINE config 4, synthetic global 1-degree ocean 360x160x32 (11.06 M unknowns,
~191.6 M non-zeros), domain-decomposed over N GPUs with the reference's Decomp2D rule (strong scaling).
A "step" is one Jacobian + one residual assembly (Ocean::computeJacobian + Ocean::computeRHS).  The same run also
times the two other parts of the path and reports them in extra keys of the same JSON line:
    "gmres": one GMRES(30) cycle (SpMV + CGS2 orthogonalisation per iteration), iterations/s
    "spmv" : the CSR SpMV alone, algorithmic GB/s
Inputs (CSR 2.3 GB, basis 2.7 GB) are larger than the 126 MB L2, so no explicit L2 flush is needed between steps.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "jacobian+rhs_assemblies_per_s"
UNIT = "assemblies/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--grid", default="360x160x32")
    ap.add_argument("--gmres-m", type=int, default=30)
    ap.add_argument("--mask", default="global", choices=["global", "none"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--skip-gmres", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (recipe of B200_PROFILING.md)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
def alg_bytes(ncell, nrows, nnz, m):
    """algorithmic HBM bytes (SURVEY.md 8d / DESIGN.md): per assembly, per SpMV, per GMRES(m) cycle"""
    jac = 932 * ncell
    rhs = 148 * ncell
    spmv = 12 * nnz + 20 * nrows
    cycle = m * spmv + (2 * m * m + 10 * m) * 8 * nrows
    return jac, rhs, spmv, cycle


def cpu_sample(threads, sample_grid=(48, 19, 12), min_seconds=8.0):
    """Times the CPU oracle (structure-faithful port of the reference's Fortran) on `threads` host threads, each
    assembling its own sample_grid block; returns cells/s for Jacobian+RHS, SpMV non-zeros/s, MGS-GMRES iteration
    cost.  This is the one place outside tests/ that executes oracle/ (checker timed as the CPU baseline)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from iemic_b200 import grids
    n, m, l = sample_grid
    cfg = grids.global_config(n, m, l)
    mask = grids.upsample_mask(grids.fixture_mask("mask_global_96x38x12"), n, m, l)
    x = grids.synthetic_state(n, m, l, mask, seed=1)
    res = [None] * threads

    def work(t):
        O = ol.Oracle(cfg, mask)
        for k, v in grids.STARTING_PARS["global"].items():
            O.set_par(k, v)
        O.rhs(x); O.jac(x)  # warm-up
        t0 = time.perf_counter(); cnt = 0
        while True:
            O.jac(x); O.rhs(x); cnt += 1
            if time.perf_counter() - t0 > min_seconds:
                break
        res[t] = (cnt, time.perf_counter() - t0)

    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    wall = time.perf_counter() - t0
    cells = sum(r[0] for r in res) * n * m * l
    secs = max(r[1] for r in res)
    return cells / secs, wall


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path (oracle port; the Fortran/Trilinos original is not
    buildable: no Fortran compiler, MPI or Trilinos in the image) on all usable host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, m, l = [int(v) for v in args.grid.split("x")]
    ncell = n * m * l
    threads = max(1, min(os.cpu_count() or 1, 32))
    per_step = max(2.0, min(10.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        cpu_sample(threads, min_seconds=1.0)
    vals = []
    t0 = time.perf_counter()
    for _ in range(max(1, min(args.steps, 5))):
        cps, _ = cpu_sample(threads, min_seconds=per_step)
        vals.append(cps / ncell)
    v = float(np.mean(vals))
    sample = "each step: %d host threads x one 48x19x12-cell block (oracle port, Jacobian+RHS), extrapolated by cell count to %s" % (threads, args.grid)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 / v, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic global 1-degree ocean %s (BASELINE config 4): Jacobian+RHS assembly per step" % args.grid,
                       "grid": args.grid, "unknowns": 6 * ncell, "mask": args.mask},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    import iemic_b200 as ib
    from iemic_b200 import grids

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            host_id = np.zeros(128, np.uint8)
            ib.capi.check(ib.lib().thcm_b200_comm_unique_id(host_id.ctypes.data))
            idbuf.copy_(torch.from_numpy(host_id))
        dist.broadcast(idbuf, 0)
        nccl_id = idbuf.cpu().numpy().copy()

    n, m, l = [int(v) for v in args.grid.split("x")]
    cfg = grids.global_config(n, m, l)
    mask = grids.synthetic_mask(n, m, l, kind=args.mask)
    G = ib.Ocean(cfg, mask, rank=rank, nranks=world, nccl_id=nccl_id)
    for k, v in grids.STARTING_PARS["global"].items():
        G.setPar(k, v)
    xg = grids.synthetic_state(n, m, l, mask)
    gid = G.global_ids()
    xloc = np.where(gid >= 0, xg[np.maximum(gid, 0)], 0.0)
    ncell_g, nrows_g = n * m * l, 6 * n * m * l
    nnz_loc = G.nnz
    nnz_t = torch.tensor([nnz_loc], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(nnz_t)
    nnz_g = int(nnz_t.item())

    dx = torch.from_numpy(xloc).cuda()
    dF = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    x_pin = torch.from_numpy(xloc[:G.nrows].copy()).pin_memory()
    F_pin = torch.empty(G.nrows, dtype=torch.float64).pin_memory()
    x_pin_np, F_pin_np = x_pin.numpy(), F_pin.numpy()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_dev():
        G.assemble(dx, dF, jacobian=True)

    def step_e2e():
        G.assemble_host(x_pin_np, F_pin_np, jacobian=True)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = G.launch_count()
    ms_dev = timed(step_dev, args.steps, args.warmup)
    launches = (G.launch_count() - l0) * args.steps // (args.steps + args.warmup)
    ms_e2e = timed(step_e2e, args.steps, max(1, args.warmup))

    # kernel-level durations (CUDA events on the launching stream, bracketing each kernel), same region re-run
    def step_seq():   # one full-grid launch per kernel, nothing concurrent: per-kernel durations for the roofline
        G.halo_exchange(dx)
        G.computeJacobian(dx)
        G.computeRHS(dx, dF)

    G.timing(True); G.timing_reset()
    timed(step_seq, args.steps, 1)
    t_jac, c_jac = G.timing_get(1)
    t_rhs, c_rhs = G.timing_get(0)
    G.timing(False)

    # SpMV alone
    ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
    y = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    ms_spmv = timed(lambda: G.applyMatrix(ones, y), max(args.steps, 20), args.warmup)
    nsp = max(args.steps, 20)

    # GMRES(m) cycle: b = J*1, x0 = 0, fixed iteration count, fused CGS2, identity preconditioner
    gm = None
    if not args.skip_gmres:
        mm = args.gmres_m
        b = y.clone()
        xs = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")

        def cycle():
            xs.zero_()
            G.solve(b, xs, m=mm, maxit=mm - 1, tol=0.0, ortho=ib.CGS2F, check_every=0)

        ksteps = max(2, min(args.steps, 5))
        ms_g = timed(cycle, ksteps, 2)
        G.timing(True); G.timing_reset()
        timed(cycle, 2, 0)
        t_orth, c_orth = G.timing_get(3)
        t_sp, c_sp = G.timing_get(2)
        G.timing(False)
        gm = (ms_g / ksteps, t_orth / max(c_orth, 1), t_sp / max(c_sp, 1))
        # SURVEY 8d "fixed preconditioner" leg (row N3): point-block Jacobi built on the device from the current Jacobian,
        # THCM block scaling, and the same GMRES cycle with the preconditioner applied (flexible: a second basis Z)
        d_binv = torch.empty(G.nrows * 6, dtype=torch.float64, device="cuda")
        d_rs = torch.empty(G.nrows, dtype=torch.float64, device="cuda")
        ms_bj = timed(lambda: G.block_jacobi(d_binv), 5, 2) / 5
        ms_sc = timed(lambda: G.recompute_scaling(d_rs, None), 5, 2) / 5
        d_ri = torch.empty_like(d_rs)
        G.reciprocal(d_rs, d_ri)

        def scale_unscale():
            G.left_scale(d_rs)
            G.left_scale(d_ri)

        ms_ls = timed(scale_unscale, 3, 1) / 6
        G.computeJacobian(dx)  # undo the rounding of scale / unscale

        def cycle_prec():
            xs.zero_()
            G.solve(b, xs, m=mm, maxit=mm - 1, tol=0.0, ortho=ib.CGS2F, prec=ib.PREC_BLOCKJACOBI, d_blockinv=d_binv, check_every=0)

        ms_gp = timed(cycle_prec, 2, 1) / 2
        prec_leg = {"block_jacobi_build_ms": ms_bj, "recompute_scaling_ms": ms_sc, "left_scale_ms": ms_ls,
                    "gmres_block_jacobi_iters_per_s": mm / (ms_gp * 1e-3), "gmres_block_jacobi_ms_per_cycle": ms_gp}
        del d_binv, d_rs, d_ri
    clocks = sampler.stop() if rank == 0 else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_kind = peaks()
    peak_total = peak * world
    jac_b, rhs_b, spmv_b, cycle_b = alg_bytes(ncell_g, nrows_g, nnz_g, args.gmres_m)
    ms_step = ms_dev / args.steps
    value = 1000.0 / ms_step
    jac_ms = t_jac / max(c_jac, 1)
    rhs_ms = t_rhs / max(c_rhs, 1)
    achieved = jac_b / world / (jac_ms * 1e-3) / 1e9  # per GPU (max-rank time approximated by rank 0's kernel time)
    traffic = None  # DRAM bytes of one jac_kernel launch from the committed ncu capture, when it is this workload
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["jac_kernel"]
        if tj["grid"] == args.grid and tj["n_gpus"] == world and args.mask == "global":
            traffic = tj["dram_bytes_per_launch"]
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "synthetic global 1-degree ocean %s (BASELINE config 4): Jacobian+RHS assembly per step" % args.grid,
                   "grid": args.grid, "unknowns": nrows_g, "nnz": nnz_g, "mask": args.mask, "decomposition": "%dx%d lon x lat" % G.block()[4:6],
                   "l2": "inputs larger than L2 (CSR %.2f GB), no flush" % (12 * nnz_g / 1e9), "gmres_m": args.gmres_m,
                       "exchange": {0: "none (one GPU)", 1: "NCCL send/recv + allreduce", 2: "peer memory over NVLink (CUDA IPC), NCCL bootstrap only"}[G.exchange_mode()]},
        "roofline": {"bound": "hbm", "kernel": "jac_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_kind": peak_kind, "algorithmic_bytes_per_launch": jac_b // world, "kernel_ms": jac_ms,
                     "rhs_kernel_ms": rhs_ms, "rhs_frac": rhs_b / world / (rhs_ms * 1e-3) / 1e9 / peak,
                     "step_frac": (jac_b + rhs_b) / (ms_step * 1e-3) / 1e9 / peak_total},
        "e2e": {"value": 1000.0 / (ms_e2e / args.steps), "unit": UNIT, "h2d_bytes_per_step": 8 * nrows_g,
                "d2h_bytes_per_step": 8 * nrows_g, "api": "thcm_b200_assemble_host: state up from pinned host memory, residual back, Jacobian device-resident like Ocean::jac_"},
        "spmv": {"gbs": spmv_b / (ms_spmv / nsp * 1e-3) / 1e9, "ms": ms_spmv / nsp, "frac": spmv_b / (ms_spmv / nsp * 1e-3) / 1e9 / peak_total,
                 "algorithmic_bytes": spmv_b},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if gm:
        line["preconditioned"] = prec_leg
        ms_cycle, orth_ms, sp_ms = gm
        line["gmres"] = {"iters_per_s": args.gmres_m / (ms_cycle * 1e-3), "ms_per_cycle": ms_cycle, "m": args.gmres_m, "ortho": "CGS2 (two reductions per step, fused normalisation)",
                         "frac": cycle_b / (ms_cycle * 1e-3) / 1e9 / peak_total, "algorithmic_bytes_per_cycle": cycle_b,
                         "orth_ms_per_step_avg": orth_ms, "spmv_ms": sp_ms}
    if not args.no_cpu_baseline:
        try:
            cps, wall = cpu_sample(1, min_seconds=10.0)
            line["cpu_baseline"] = {"value": cps / ncell_g, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "oracle port, Jacobian+RHS on one 48x19x12-cell block for %.0f s on 1 core, extrapolated by cell count" % wall}
        except Exception as e:  # the checker is missing: report, do not fake
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": "unavailable: %s" % e}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
