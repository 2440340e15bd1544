"""2-GPU debug: where does the fused assembly differ from the stand-alone Jacobian kernel on each rank?"""
import os, sys
import numpy as np, torch, torch.distributed as dist
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases, iemic_b200 as ib
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    hid = np.zeros(128, np.uint8); ib.capi.check(ib.lib().thcm_b200_comm_unique_id(hid.ctypes.data)); idbuf.copy_(torch.from_numpy(hid))
dist.broadcast(idbuf, 0)
name = sys.argv[1] if len(sys.argv) > 1 else "rand_periodic"
cfg, mask, pars = cases.case(name)
G = ib.Ocean(cfg, mask, rank=rank, nranks=world, nccl_id=idbuf.cpu().numpy().copy())
for k, v in pars.items(): G.setPar(k, v)
xg = cases.state(cfg, mask, seed=21)
gid = G.global_ids(); own = gid[:G.nrows]
vg, bg = G.computeJacobian_host(xg[own])
Fg = G.computeRHS_host(xg[own])
dxs = torch.full((G.nvec,), 7.0, dtype=torch.float64, device="cuda"); dxs[:G.nrows] = torch.from_numpy(xg[own]).cuda()
dFs = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
for trial in range(3):
    G.set_values_host(np.full(G.nnz, 777.0))
    G.assemble(dxs, dFs, jacobian=True); G.sync()
    v2 = G.values_host()[0]
    rp, col = G.graph()
    rows = np.repeat(np.arange(G.nrows), np.diff(rp))
    bad = np.flatnonzero(v2 != vg)
    i0, j0, nloc, mloc, pn, pm = G.block()
    msg = "rank %d trial %d block i0 %d j0 %d nloc %d mloc %d: bad %d of %d, F equal %s" % (rank, trial, i0, j0, nloc, mloc, bad.size, vg.size, np.array_equal(dFs.cpu().numpy()[:G.nrows], Fg))
    for q in bad[:12]:
        r = rows[q]; c = r // 6
        msg += "\n   nz %d row %d var %d li %d lj %d k %d pos %d/%d fused %r alone %r" % (q, r, r % 6, c % nloc, (c // nloc) % mloc, c // (nloc * mloc), q - rp[r], rp[r + 1] - rp[r], v2[q], vg[q])
    print(msg, flush=True)
dist.barrier()
dist.destroy_process_group()
