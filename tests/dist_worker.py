"""Worker of tests/test_dist_cpu.py::test_gloo_world2_halo_exchange (launched by torch.distributed.run, gloo, CPU).

Each rank builds its block of the decomposition, fills only its OWNED part of the state, exchanges halos with the plan
(send/recv lists of thcm::Host::halo_plan -- the same lists the NCCL path packs with), assembles residual and Jacobian with
the product's row code and checks them against the single-domain oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import hostcheck_lib as hl  # noqa: E402
import oracle_lib as ol  # noqa: E402


def exchange(H, x, rank):
    """x: stored vector with owned part filled; fills the ring through the plan"""
    plan = H.halo_plan()
    x6 = x.reshape(-1, 6)
    reqs, bufs = [], []
    for peer, send, recv in plan:
        if peer == rank:
            x6[recv] = x6[send]
            continue
        sb = torch.from_numpy(np.ascontiguousarray(x6[send]))
        rb = torch.empty((recv.size, 6), dtype=torch.float64)
        reqs.append(dist.isend(sb, peer))
        reqs.append(dist.irecv(rb, peer))
        bufs.append((recv, rb, sb))
    for r in reqs:
        r.wait()
    for recv, rb, _ in bufs:
        x6[recv] = rb.numpy()


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    for name in ("rand_periodic", "rand_box", "global_half"):
        cfg, mask, pars = cases.case(name)
        H = hl.HostCheck(cfg, mask, rank=rank, nranks=world)
        for k, v in pars.items():
            H.set_par(k, v)
        xg = cases.state(cfg, mask, seed=12)
        gid = H.gid()
        x = np.zeros(H.nvec)
        x[:H.ndim] = xg[gid[:H.ndim]]
        exchange(H, x, rank)
        used = gid >= 0
        assert np.array_equal(x[used], xg[gid[used]]), "halo exchange delivered wrong values"
        O = ol.Oracle(cfg, mask)
        for k, v in pars.items():
            O.set_par(k, v)
        assert np.array_equal(H.rhs(x), O.rhs(xg)[gid[:H.ndim]])
        # global dot through an allreduce (the GMRES reduction path)
        part = torch.tensor([float(np.dot(x[:H.ndim], x[:H.ndim]))], dtype=torch.float64)
        dist.all_reduce(part)
        assert abs(part.item() - float(np.dot(xg, xg))) <= 1e-12 * float(np.dot(xg, xg))
    print("rank %d ok" % rank, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
