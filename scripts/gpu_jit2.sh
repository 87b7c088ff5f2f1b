mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_jit.py -x -q > gpurun_out/pytest_jit.log 2>&1; echo "pytest jit rc=$?"; tail -15 gpurun_out/pytest_jit.log
B200SV_JIT=always B200SV_JIT_DBUF=2 timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_all_jit_always.log 2>&1; echo "pytest all(jit always) rc=$?"; tail -15 gpurun_out/pytest_all_jit_always.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_jit.csv python scripts/jit30.py 30 4:11:1 > gpurun_out/ncu_list_jit.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pass -s 34 -c 6 -f -o gpurun_out/jit_full python scripts/jit30.py 30 4:11:1 > gpurun_out/ncu_full_jit.log 2>&1; echo "ncu full rc=$?"
