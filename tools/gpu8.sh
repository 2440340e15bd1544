# 8-GPU evidence: multi-GPU parity (2/4/8 ranks x peer/NCCL) with its log, then the bench at N = 8, 4, 2
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dist.py -m gpu -v > gpurun_out/dist_parity_8gpu.log 2>&1; tail -9 gpurun_out/dist_parity_8gpu.log
for n in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_r10_n$n.json 2> gpurun_out/bench_r10_n$n.err
tail -c 2500 gpurun_out/bench_r10_n$n.json | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print(d['n_gpus'], d['value'], d['e2e']['value'], d['gmres']['iters_per_s'], d['spmv']['ms'], d['parity']['status'], d['parity'].get('rel_dev'))
except Exception as e: print('parse failed', e)
"
done
