#!/bin/bash
# round 2, GPU call B (1 GPU): config-size parity first (verbose), then the rest of the suite, per-pass log, bench
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt
nproc >> gpurun_out/gpu.txt; free -g | head -2 >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -s > gpurun_out/pytest_configs.log 2>&1; echo "configs rc=$?"
grep "^\[parity\]\|passed\|failed\|Error\|assert" gpurun_out/pytest_configs.log | cut -c1-260 | head -60
timeout 1500 python -m pytest tests -m gpu -q -s --deselect tests/test_gpu_configs.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Error" gpurun_out/pytest_gpu.log | cut -c1-260 | head -40
B200SV_JIT_CACHE=off B200SV_PASS_LOG=1 timeout 600 python scripts/jit30.py 30 4:11:1:3 > gpurun_out/passlog.txt 2>&1
grep "b200sv pass" gpurun_out/passlog.txt | tail -18
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "cold_start", "extra", "jit"):
        print(k, json.dumps(j.get(k))[:1500])
except Exception as e:
    print("bench parse failed", e)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:k_pass -c 18 -f -o /tmp/pass_full python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/pass_full.ncu-rep --page details --csv > gpurun_out/pass_ncu_details.csv 2>/dev/null
ncu -i /tmp/pass_full.ncu-rep --page raw --csv > gpurun_out/pass_ncu_raw.csv 2>/dev/null
ls -la gpurun_out /tmp/pass_full.ncu-rep | tail -20
du -sh gpurun_out
