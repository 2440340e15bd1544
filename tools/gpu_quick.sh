# quick GPU check of the assembly kernels: parity tests, timings of the 1-degree grid, optional ncu capture (PROF=tag)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or rhs_parity or jacobian_parity or assemble_host or full_size" > gpurun_out/quick_tests.log 2>&1
tail -3 gpurun_out/quick_tests.log
python tools/jac_variants.py 2>&1 | tail -2 | tee gpurun_out/quick_times.txt
MASK=none python tools/jac_variants.py 2>&1 | tail -1 | tee -a gpurun_out/quick_times.txt
if [ -n "$PROF" ]; then
python tools/prof_asm.py > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"asm_kernel" -s 3 -c 3 -o gpurun_out/prof_$PROF python tools/prof_asm.py > gpurun_out/ncu_$PROF.log 2>&1
tail -2 gpurun_out/ncu_$PROF.log
fi
