#!/bin/bash
# round 2, GPU call O (1 GPU): rewritten k_bilinear_batch: gradient / expval parity, then C2 / C4 timings
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_jit.py tests/test_gpu_plugin.py -m gpu -q -s --timeout 240 -p no:cacheprovider -k "grad or expval or c2 or c4 or observ or golden or var" > gpurun_out/pytest_grad.log 2>&1; echo "grad tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_grad.log | cut -c1-300 | head; grep "^\[parity\] C[24]" gpurun_out/pytest_grad.log | cut -c1-200
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "e2e", j["e2e"]["value"])
    print("extra", json.dumps(j.get("extra")))
except Exception as ex:
    print("bench parse failed", ex)
PY
