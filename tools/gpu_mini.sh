# after a change of thcm_kernels.cu: parity of the assembly kernels incl. the 1-degree golden, bench line, one full capture of the fused kernel
tag=${1:-r11}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fortran_abi.py -m gpu -x -q -k "fused or rhs_parity or jacobian_parity or assemble_host or golden or parameter or fortran" > gpurun_out/gputests_$tag.log 2>&1; tail -2 gpurun_out/gputests_$tag.log
timeout 300 python bench.py > gpurun_out/bench_${tag}_n1.json 2> gpurun_out/bench_${tag}_n1.err; tail -c 300 gpurun_out/bench_${tag}_n1.json
timeout 100 python tools/prof_asm.py > gpurun_out/plain_prof.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"asm_kernel" -s 6 -c 3 -o gpurun_out/prof_$tag python tools/prof_asm.py > gpurun_out/ncu_prof_$tag.log 2>&1
tail -1 gpurun_out/ncu_prof_$tag.log
