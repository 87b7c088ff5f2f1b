#!/bin/bash
# round 2, final GPU call (1 GPU): full suite (-x, the no-alloc test last), smoke(), default bench
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_gpu.log | cut -c1-300 | head -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "ms", j["ms_per_step"], "frac", j["roofline"]["frac"], "launches", j["gpu_launches"], "clocks", j["clocks"])
    e = j["e2e"]
    print("e2e", e["value"], {k: v for k, v in e.items() if k.endswith("_ms") and not isinstance(v, dict)})
    print("cpu", j["cpu_baseline"]["value"] if j.get("cpu_baseline") else None, "second process", json.dumps(j.get("cold_start", {}).get("second_process")))
    print("extra", json.dumps({k: {a: round(b, 1) for a, b in v.items() if a.endswith("_ms")} for k, v in (j.get("extra") or {}).items()}))
except Exception as ex:
    print("bench parse failed", ex)
PY
