#!/bin/bash
# round 2, GPU call K (1 GPU): kept allocations + streaming host copy: targeted tests, bench, then the whole suite -x
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gpu_zz_reference_programs.py tests/test_gpu_plugin.py -m gpu -q --timeout 240 -p no:cacheprovider > gpurun_out/pytest_keep.log 2>&1; echo "keep tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_keep.log | cut -c1-300 | head
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "ms", j["ms_per_step"], "frac", j["roofline"]["frac"])
    e = j["e2e"]
    print("e2e", e["value"], "A", e["run_a_probs_ms"], "B", e["run_b_sample_ms"], "split", json.dumps(e.get("host_split_ms", {}).get("run_a")), json.dumps(e.get("host_split_ms", {}).get("run_b")))
except Exception as ex:
    print("bench parse failed", ex)
PY
B200SV_NO_STREAM_COPY=1 timeout 600 python bench.py --no-extra --no-cpu-baseline > gpurun_out/bench_nostream.json 2>/dev/null; python - <<'PY'
import json
try:
    e = json.loads(open("gpurun_out/bench_nostream.json").read().strip().splitlines()[-1])["e2e"]
    print("memcpy instead of streaming copy: A", e["run_a_probs_ms"], "split", json.dumps(e["host_split_ms"]["run_a"]))
except Exception as ex:
    print("parse failed", ex)
PY
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_gpu.log | cut -c1-300 | head -20
