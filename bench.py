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
    ap.add_argument("--write-parity-ref", action="store_true", help="one GPU only: record J1_norm2 / gmres_hist of this run as the committed reference")
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


def probe_reference_toolchain():
    """can the reference's own Fortran / Trilinos path be built on THIS box?  Probed at run time."""
    import shutil
    have = {t: bool(shutil.which(t)) for t in ("gfortran", "mpirun", "mpicxx", "cmake")}
    epetra = False
    for root in ("/usr/include", "/usr/local/include", "/opt"):
        if not os.path.isdir(root):
            continue
        for dp, dn, fn in os.walk(root):
            if dp.count(os.sep) > 6:
                dn[:] = []
            if "Epetra_CrsMatrix.h" in fn:
                epetra = True
                break
        if epetra:
            break
    have["Epetra_CrsMatrix.h"] = epetra
    if have["gfortran"] and have["mpirun"] and epetra:
        return "reference CPU path: toolchain present (gfortran, MPI, Trilinos) -- not built by this harness", have
    missing = [k for k in ("gfortran", "mpirun", "Epetra_CrsMatrix.h") if not have[k]]
    return "reference CPU path: not buildable on this box (missing: %s)" % ", ".join(missing), have


def host_threads():
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return max(1, min(n, 64))


class CpuArm:
    """The reference's CPU algorithm on the STATED grid: the n x m x l problem is cut into `threads` blocks by the reference's own
    Decomp2D rule, every block is padded with the two ghost layers the reference uses (TRIOS_Domain.C:182-189), and one instance of
    the CPU oracle (oracle/: slot-for-slot port of the Fortran, including the row copy into the static graph of THCM.C:1074-1155)
    assembles Jacobian + residual of its block on its own thread.  One step = the whole grid, wall clock, no extrapolation.
    This is the one place outside tests/ that executes oracle/ (the checker, timed as the CPU baseline)."""

    GHOST = 2

    def __init__(self, grid, mask_kind, threads):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        from iemic_b200 import grids
        self.ol, self.grids = ol, grids
        n, m, l = grid
        self.grid, self.threads = grid, threads
        self.cfg = grids.global_config(n, m, l)
        self.mask = grids.synthetic_mask(n, m, l, kind=mask_kind)
        self.x = grids.synthetic_state(n, m, l, self.mask).reshape(l, m, n, 6)
        self.npN, self.npM = grids.decomp2d(n, m, threads)
        self.blocks = []
        for pm in range(self.npM):
            for pn in range(self.npN):
                self.blocks.append(self._block(pn, pm))
        self.cells_with_ghosts = sum(b[0]["n"] * b[0]["m"] * l for b in self.blocks)
        self.oracles = None

    @staticmethod
    def _split(n, parts, p):
        base, rem = divmod(n, parts)
        lo = p * base + min(p, rem)
        return lo, lo + base + (1 if p < rem else 0)

    def _block(self, pn, pm):
        n, m, l = self.grid
        g = self.GHOST
        i0, i1 = self._split(n, self.npN, pn)
        j0, j1 = self._split(m, self.npM, pm)
        perio = bool(self.cfg["periodic"])
        gw = g if (self.npN > 1 and (perio or pn > 0)) else 0
        ge = g if (self.npN > 1 and (perio or pn < self.npN - 1)) else 0
        gs = g if pm > 0 else 0
        gn = g if pm < self.npM - 1 else 0
        ii = np.arange(i0 - gw, i1 + ge) % n                       # owned + ghost columns (wrapped)
        jj = np.arange(j0 - gs, j1 + gn)
        nb, mb = ii.size, jj.size
        sub = np.ones((l + 2, mb + 2, nb + 2), np.int32)
        block_periodic = perio and self.npN == 1
        ring_j = np.clip(np.concatenate(([jj[0] - 1], jj, [jj[-1] + 1])), -1, m) + 1      # indices into the (m + 2) mask
        if block_periodic:
            ring_i = np.arange(n + 2)                                                      # the global mask's own ring (PERIO markers)
        else:
            ring_i = np.concatenate(([ii[0] - 1], ii, [ii[-1] + 1])) % n + 1
        full = self.mask
        sub[:, :, :] = full[np.ix_(np.arange(l + 2), ring_j, ring_i)]
        if not block_periodic:                                       # a cut side is a wall for the sub-domain (usrc.F90:96-107)
            sub[:, :, 0] = 1; sub[:, :, -1] = 1
        sub[:, 0, :] = np.where(jj[0] == 0, sub[:, 0, :], 1); sub[:, -1, :] = np.where(jj[-1] == m - 1, sub[:, -1, :], 1)
        cfg = dict(self.cfg)
        dx = (self.cfg["xmax"] - self.cfg["xmin"]) / n
        dy = (self.cfg["ymax"] - self.cfg["ymin"]) / m
        cfg.update(n=nb, m=mb, xmin=self.cfg["xmin"] + (i0 - gw) * dx, xmax=self.cfg["xmin"] + (i1 + ge) * dx,
                   ymin=self.cfg["ymin"] + (j0 - gs) * dy, ymax=self.cfg["ymin"] + (j1 + gn) * dy, periodic=int(block_periodic))
        xs = np.ascontiguousarray(self.x[np.ix_(np.arange(l), jj, ii, np.arange(6))]).ravel()
        return cfg, sub, xs

    def bytes_needed(self):
        return int(self.cells_with_ghosts * 21e3)                     # An + Al (2 x 972 doubles per cell) + CSR + fields

    def setup(self):
        """one oracle per block, created on its own thread (Al is built here: parameter-change cost, not assembly cost)"""
        pars = self.grids.STARTING_PARS["global"]
        self.oracles = [None] * len(self.blocks)

        def mk(t):
            cfg, sub, xs = self.blocks[t]
            O = self.ol.Oracle(cfg, sub)
            for k, v in pars.items():
                O.set_par(k, v)
            self.oracles[t] = O

        self._run(mk)

    def _run(self, fn):
        th = [threading.Thread(target=fn, args=(t,)) for t in range(len(self.blocks))]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        return time.perf_counter() - t0

    def step(self):
        """one Jacobian + residual assembly of the whole grid; returns wall seconds"""
        def work(t):
            O, xs = self.oracles[t], self.blocks[t][2]
            O.jac(xs)
            O.rhs(xs)

        return self._run(work)

    def describe(self):
        n, m, l = self.grid
        return ("oracle port (kind=port), %dx%dx%d cut by Decomp2D into %dx%d blocks (+%d ghost layers: %d cells assembled for %d owned), one block per host "
                "thread, %d threads; each step assembles the WHOLE grid (Jacobian incl. the row copy into the static graph + residual), wall clock"
                % (n, m, l, self.npN, self.npM, self.GHOST, self.cells_with_ghosts, n * m * l, self.threads))


def cpu_arm_for(grid, mask_kind, threads):
    """CpuArm sized to the memory of this box: fewer threads (= fewer, larger blocks) never needs more memory, so only refuse when even
    the bare grid does not fit"""
    arm = CpuArm(grid, mask_kind, threads)
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = None
    if avail is not None and arm.bytes_needed() > 0.9 * avail:
        raise MemoryError("CPU oracle needs about %.0f GB for this grid, %.0f GB available" % (arm.bytes_needed() / 1e9, avail / 1e9))
    return arm


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the path (oracle port; the Fortran/Trilinos original is not
    buildable: no Fortran compiler, MPI or Trilinos in the image) on all usable host threads, on the stated grid."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, m, l = [int(v) for v in args.grid.split("x")]
    ncell = n * m * l
    threads = host_threads()
    probe, _ = probe_reference_toolchain()
    t_all = time.perf_counter()
    arm = cpu_arm_for((n, m, l), args.mask, threads)
    t0 = time.perf_counter()
    arm.setup()
    setup_s = time.perf_counter() - t0
    budget = 170.0                                                     # seconds of timed + warm-up steps
    first = arm.step()                                                 # warm-up step, also sizes the run
    warm = max(0, min(args.warmup, int(0.25 * budget / max(first, 1e-3))) - 1)
    for _ in range(warm):
        arm.step()
    steps = max(1, min(args.steps, int(0.75 * budget / max(first, 1e-3))))
    secs = [arm.step() for _ in range(steps)]
    v = steps / sum(secs)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "steps_run": steps, "warmup_run": warm + 1, "ms_per_step": 1000.0 / v, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic global 1-degree ocean %s (BASELINE config 4): Jacobian+RHS assembly per step" % args.grid,
                       "grid": args.grid, "unknowns": 6 * ncell, "mask": args.mask},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": arm.describe()},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "same_config": True, "extrapolated": False, "step_seconds": {"min": min(secs), "median": float(np.median(secs)), "max": max(secs)},
            "setup_seconds": setup_s, "reference_toolchain": probe,
            "note": "the port omits nothing of THCM::evaluate's Jacobian branch except Epetra's FillComplete; ghost cells are assembled and discarded like the reference does",
            "gpu_launches": 0, "wall_s": time.perf_counter() - t_all}
    print(json.dumps(line))


PARITY_REF = os.path.join(ROOT, "tests", "golden", "bench_parity_ref.json")


def check_parity(p, args, world):
    """compares the run's decomposition-invariant numbers with the committed one-GPU values (tests/golden/bench_parity_ref.json:
    F_norm2 there is the CPU ORACLE's value from tests/golden/global1deg_golden.npz; J1 and the GMRES history were taken on one B200)"""
    try:
        ref = json.load(open(PARITY_REF))["%s/%s/m%d" % (args.grid, args.mask, args.gmres_m)]
    except Exception:
        return {"status": "no reference for this workload", "tol": None}
    out = {"status": "ok", "tol": {"norms": 1e-8, "gmres_hist": 1e-6}, "rel_dev": {}}
    for k, tol in (("F_norm2", 1e-8), ("J1_norm2", 1e-8)):
        d = abs(p[k] - ref[k]) / abs(ref[k])
        out["rel_dev"][k] = d
        if not d <= tol:
            out["status"] = "FAILED"
    if "gmres_hist" in p and "gmres_hist" in ref:
        a, b = np.array(p["gmres_hist"]), np.array(ref["gmres_hist"])
        d = float(np.max(np.abs(a - b) / np.abs(b))) if a.shape == b.shape else float("inf")
        out["rel_dev"]["gmres_hist"] = d
        if not d <= 1e-6:
            out["status"] = "FAILED"
    return out


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

    # ---- parity gate (the reference's I5 idiom, src/tests/test_matrix.C:148-192): decomposition-invariant numbers of THIS run's
    # distributed path, compared with the committed one-GPU values before anything is timed
    def gsum(t):
        if world > 1:
            dist.all_reduce(t)
        return float(t.item())

    ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
    y = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    G.assemble(dx, dF, jacobian=True)
    G.sync()
    F_norm = gsum((dF[:G.nrows] ** 2).sum()) ** 0.5
    G.applyMatrix(ones, y)
    G.sync()
    J1_norm = gsum((y[:G.nrows] ** 2).sum()) ** 0.5
    parity = {"F_norm2": F_norm, "J1_norm2": J1_norm}
    if not args.skip_gmres:
        xs0 = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        rc_p, hist_p, it_p = G.solve(y.clone(), xs0, m=args.gmres_m, maxit=args.gmres_m - 1, tol=0.0, ortho=ib.CGS2F, check_every=0)
        parity["gmres_hist"] = [float(v) for v in hist_p[:args.gmres_m][::5]] + [float(hist_p[args.gmres_m - 1])]
        del xs0
    if args.write_parity_ref and world == 1 and rank == 0:
        key = "%s/%s/m%d" % (args.grid, args.mask, args.gmres_m)
        ref_all = json.load(open(PARITY_REF)) if os.path.exists(PARITY_REF) else {}
        entry = dict(parity, source="J1_norm2 and gmres_hist: bench.py --write-parity-ref on one B200")
        gold = os.path.join(ROOT, "tests", "golden", "global1deg_golden.npz")
        if args.grid == "360x160x32" and args.mask == "global" and os.path.exists(gold):
            oracle_norm = float(np.load(gold)["norms"][0])
            entry["F_norm2_this_run"] = entry["F_norm2"]
            entry["F_norm2"] = oracle_norm
            entry["source"] += "; F_norm2: CPU oracle (tests/golden/global1deg_golden.npz)"
        ref_all[key] = entry
        json.dump(ref_all, open(PARITY_REF, "w"), indent=1)
    parity.update(check_parity(parity, args, world))
    if parity["status"] == "FAILED":
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": None, "unit": UNIT, "n_gpus": world, "parity": parity,
                              "error": "parity gate failed: the distributed path does not reproduce the committed one-GPU numbers"}))
        if world > 1:
            dist.destroy_process_group()
        raise SystemExit(3)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = G.launch_count()
    ms_dev = timed(step_dev, args.steps, args.warmup)
    launches = (G.launch_count() - l0) * args.steps // (args.steps + args.warmup)
    reps_ms = [ms_dev / args.steps] + [timed(step_dev, args.steps, 0) / args.steps for _ in range(4)]   # run-to-run noise of the same loop
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
    timed(step_dev, args.steps, 1)        # the step's own kernel: Jacobian + residual fused in one launch
    t_asm, c_asm = G.timing_get(4)
    G.timing(False)

    # SpMV alone
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
    # row N1: what a parameter change costs on top of the next assembly (Continuation::computeDFDPar changes one twice per Newton step)
    kk = [0]

    def step_setpar():
        kk[0] += 1
        G.setPar("COMB", 0.1 + 1.0e-9 * kk[0])
        G.assemble(dx, dF, jacobian=True)

    ms_sp = timed(step_setpar, args.steps, 2) / args.steps
    ms_plain = timed(step_dev, args.steps, 1) / args.steps
    G.setPar("COMB", grids.STARTING_PARS["global"]["COMB"])
    set_par_ms = ms_sp - ms_plain
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
    asm_ms = t_asm / max(c_asm, 1)
    # dominant kernel of the step: asm_kernel<JAC,RHS> (one launch per step): 1080 B per cell
    achieved = (jac_b + rhs_b) / world / (asm_ms * 1e-3) / 1e9  # per GPU (max-rank time approximated by rank 0's kernel time)
    traffic, traffic_src = None, None  # DRAM bytes of one launch from the committed ncu capture -- only if taken from THIS kernel source
    try:
        import hashlib
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["asm_kernel"]
        sha = hashlib.sha256(open(os.path.join(ROOT, "iemic_b200", "csrc", "thcm_kernels.cu"), "rb").read()).hexdigest()[:16]
        if tj["grid"] == args.grid and tj["n_gpus"] == world and args.mask == "global" and tj.get("source_sha") == sha:
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj.get("capture")
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
        "roofline": {"bound": "hbm", "kernel": "asm_kernel<JAC,RHS> (Jacobian + residual, one launch per step)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_kind": peak_kind,
                     "algorithmic_bytes_per_launch": (jac_b + rhs_b) // world, "kernel_ms": asm_ms,
                     "jacobian_alone_ms": jac_ms, "jacobian_alone_frac": jac_b / world / (jac_ms * 1e-3) / 1e9 / peak,
                     "residual_alone_ms": rhs_ms, "residual_alone_frac": rhs_b / world / (rhs_ms * 1e-3) / 1e9 / peak,
                     "step_frac": (jac_b + rhs_b) / (ms_step * 1e-3) / 1e9 / peak_total},
        "repeat": {"loops": len(reps_ms), "ms_per_step_min": min(reps_ms), "ms_per_step_median": float(np.median(reps_ms)), "ms_per_step_max": max(reps_ms)},
        "parity": parity,
        "e2e": {"value": 1000.0 / (ms_e2e / args.steps), "unit": UNIT, "h2d_bytes_per_step": 8 * nrows_g,
                "d2h_bytes_per_step": 8 * nrows_g, "api": "thcm_b200_assemble_host: state up from pinned host memory, residual back, Jacobian device-resident like Ocean::jac_"},
        "spmv": {"gbs": spmv_b / (ms_spmv / nsp * 1e-3) / 1e9, "ms": ms_spmv / nsp, "frac": spmv_b / (ms_spmv / nsp * 1e-3) / 1e9 / peak_total,
                 "algorithmic_bytes": spmv_b},
        "set_par_ms": set_par_ms,
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if gm:
        line["preconditioned"] = prec_leg
        ms_cycle, orth_ms, sp_ms = gm
        line["gmres"] = {"iters_per_s": args.gmres_m / (ms_cycle * 1e-3), "ms_per_cycle": ms_cycle, "m": args.gmres_m, "ortho": "CGS2 (two reductions per step, fused normalisation)",
                         "frac": cycle_b / (ms_cycle * 1e-3) / 1e9 / peak_total, "algorithmic_bytes_per_cycle": cycle_b,
                         "orth_ms_per_step_avg": orth_ms, "spmv_ms": sp_ms}
    line["reference_toolchain"] = probe_reference_toolchain()[0]
    if not args.no_cpu_baseline and world == 1:
        try:
            threads = host_threads()
            arm = cpu_arm_for((n, m, l), args.mask, threads)
            arm.setup()
            sec = arm.step()
            line["cpu_baseline"] = {"value": 1.0 / sec, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "ONE whole-grid assembly (%.1f s wall): %s" % (sec, arm.describe())}
        except Exception as e:  # the checker is missing or does not fit: report, do not fake
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": "unavailable: %s" % e}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
