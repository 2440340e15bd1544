# round-end evidence on one B200: smoke, the whole GPU suite, the bench line, the ncu launch list of the bench command and one
# full capture of the hot kernels.  usage: gpurun --timeout 1200 -- 'bash tools/gpu_final.sh r10'
tag=${1:-r10}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; tail -1 gpurun_out/smoke_$tag.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gputests_$tag.log 2>&1; tail -2 gpurun_out/gputests_$tag.log
timeout 300 python bench.py > gpurun_out/bench_${tag}_n1.json 2> gpurun_out/bench_${tag}_n1.err; tail -c 600 gpurun_out/bench_${tag}_n1.json
timeout 200 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_launches.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_$tag.log 2>&1
timeout 100 python tools/profile_step.py 360x160x32 1 > gpurun_out/plain_prof.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"asm_kernel|spmv_tile|multi_axpy_dot|multi_dot_kernel" -c 10 -o gpurun_out/prof_$tag python tools/profile_step.py 360x160x32 1 > gpurun_out/ncu_prof_$tag.log 2>&1
tail -1 gpurun_out/ncu_prof_$tag.log
