"""Latency breakdown of the exchange steps on N GPUs (run under torchrun, one rank per GPU).

Prints, on rank 0, device-timed averages (max over ranks) of: the halo exchange alone, a 31-double NCCL allreduce,
the sequential and the overlapped assembly, SpMV (halo included), and one GMRES(30) cycle per orthogonalisation mode.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import iemic_b200 as ib  # noqa: E402
from iemic_b200 import grids  # noqa: E402


def main():
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        hid = np.zeros(128, np.uint8)
        ib.capi.check(ib.lib().thcm_b200_comm_unique_id(hid.ctypes.data))
        idbuf.copy_(torch.from_numpy(hid))
    dist.broadcast(idbuf, 0)
    n, m, l = 360, 160, 32
    cfg = grids.global_config(n, m, l)
    mask = grids.synthetic_mask(n, m, l, kind="global")
    G = ib.Ocean(cfg, mask, rank=rank, nranks=world, nccl_id=idbuf.cpu().numpy().copy())
    for k, v in grids.STARTING_PARS["global"].items():
        G.setPar(k, v)
    xg = grids.synthetic_state(n, m, l, mask)
    gid = G.global_ids()
    dx = torch.from_numpy(np.where(gid >= 0, xg[np.maximum(gid, 0)], 0.0)).cuda()
    dF = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    y = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    small = torch.zeros(31, dtype=torch.float64, device="cuda")

    def timed(fn, reps, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def seq():
        G.halo_exchange(dx); G.computeJacobian(dx); G.computeRHS(dx, dF)

    out = {"n_gpus": world, "exchange_mode": G.exchange_mode()}
    out["halo_ms"] = timed(lambda: G.halo_exchange(dx), 200)
    out["allreduce31_ms"] = timed(lambda: dist.all_reduce(small), 200)
    out["jac_only_ms"] = timed(lambda: G.computeJacobian(dx), 50)
    out["rhs_only_ms"] = timed(lambda: G.computeRHS(dx, dF), 50)
    out["assemble_seq_ms"] = timed(seq, 50)
    for plan in (0, 1, 2):
        os.environ["IEMIC_ASM_PLAN"] = str(plan)
        out["assemble_plan%d_ms" % plan] = timed(lambda: G.assemble(dx, dF, jacobian=True), 50)
    os.environ.pop("IEMIC_ASM_PLAN")
    ones = torch.ones(G.nvec, dtype=torch.float64, device="cuda")
    out["spmv_ms"] = timed(lambda: G.applyMatrix(ones, y), 100)
    b = y.clone()
    xs = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
    for name, mode in (("cgs2", ib.CGS2), ("cgs2f", ib.CGS2F), ("mgs", ib.MGS)):
        def cycle():
            xs.zero_()
            G.solve(b, xs, m=30, maxit=29, tol=0.0, ortho=mode, check_every=0)
        out["gmres30_%s_ms" % name] = timed(cycle, 3, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    G.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
