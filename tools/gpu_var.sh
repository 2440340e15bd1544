# times the assembly kernels of the default library and of every variant under iemic_b200/lib/variants
mkdir -p gpurun_out
python tools/jac_variants.py 2>&1 | tail -1 | tee gpurun_out/var_times.txt
for f in iemic_b200/lib/variants/*.so; do IEMIC_B200_LIB=$PWD/$f python tools/jac_variants.py 2>&1 | tail -1 | tee -a gpurun_out/var_times.txt; done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or rhs_parity or jacobian_parity" 2>&1 | tail -2
