"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): NCCL halo exchange + allreduces against the single-domain oracle."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("p2p", ["peer", "nccl"])
@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_multi_gpu_parity(nproc, p2p):
    """Both exchange paths: peer memory over NVLink (the default when CUDA IPC is available) and NCCL."""
    if _ngpu() < nproc:
        pytest.skip("needs %d GPUs" % nproc)
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, IEMIC_P2P="1" if p2p == "peer" else "0", IEMIC_DEBUG="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + nproc + (10 if p2p == "nccl" else 0)), os.path.join(here, "dist_gpu_worker.py")]
    # own process group + hard kill: a rank that fails while its peers wait in a collective must not leave workers spinning on the GPUs
    p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env, start_new_session=True)
    try:
        out, err = p.communicate(timeout=300)
    except subprocess.TimeoutExpired:
        import signal
        os.killpg(p.pid, signal.SIGKILL)
        out, err = p.communicate()
        raise AssertionError("multi-GPU worker timed out\n" + out[-2000:] + err[-3000:])
    log = os.path.join(os.path.dirname(here), "gpurun_out")
    if os.path.isdir(log):
        with open(os.path.join(log, "dist_worker_%d_%s.log" % (nproc, p2p)), "w") as f:
            f.write(out + "\n---- stderr ----\n" + err)
    r = subprocess.CompletedProcess(cmd, p.returncode, out, err)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    for k in range(nproc):
        assert "rank %d ok" % k in r.stdout
    if p2p == "nccl":
        assert "peer-memory exchange on" not in r.stderr
