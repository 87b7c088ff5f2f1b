mkdir -p gpurun_out
B200SV_JIT_VERBOSE=1 timeout 1200 python -m pytest tests/test_gpu_jit.py -x -q > gpurun_out/pytest_jit.log 2>&1; echo "pytest rc=$?"
grep -v "^\[b200sv jit\] k_pass" gpurun_out/pytest_jit.log | tail -30
timeout 900 python scripts/jit30.py 30 ${CFGS:-4:11:1} > gpurun_out/jit30.log 2>&1; echo "jit30 rc=$?"; cat gpurun_out/jit30.log
