# Runs on the GPU box (gpurun): the reference Continuation.H through the seven legs of intt_2dmoc on the CUDA path, then the
# GMRES tests; everything it writes goes to gpurun_out/.  usage: gpurun --timeout 420 -- "bash tools/gpu_chain_check.sh"
mkdir -p gpurun_out/chain
( time timeout 200 env IEMIC_COMPAT_VERBOSE=1 tests/_build/continuation_dropin chain gpurun_out/chain ) > gpurun_out/cd_chain2.log 2>&1; echo "chain rc=$?" > gpurun_out/chain_rc.txt
grep -a "^leg \|drop-in\|real" gpurun_out/cd_chain2.log > gpurun_out/chain_summary.txt
grep -ac "GMRES rc = 2" gpurun_out/cd_chain2.log >> gpurun_out/chain_summary.txt
gzip -f gpurun_out/cd_chain2.log
timeout 170 python -m pytest tests/test_gpu_gmres_guard.py tests/test_gpu_parity.py tests/test_gpu_scaling.py -m gpu -x -q -k "gmres or guard or solve" > gpurun_out/guard_tests.log 2>&1
tail -5 gpurun_out/guard_tests.log; cat gpurun_out/chain_summary.txt
