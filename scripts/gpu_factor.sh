mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_jit.py -x -q > gpurun_out/pytest_jit.log 2>&1; echo "pytest jit rc=$?"; tail -8 gpurun_out/pytest_jit.log
B200SV_JIT_FACTOR=0 timeout 600 python scripts/jit30.py 30 4:11:1:3 2>&1 | grep -v waited | tail -3
B200SV_JIT_FACTOR=1 timeout 600 python scripts/jit30.py 30 4:11:1:3 2>&1 | grep -v waited | tail -3
