#!/bin/bash
# round 2, GPU call M (1 GPU): conflict-free |psi|^2 permutation kernel + coalesced sampler preparation
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q --timeout 240 -p no:cacheprovider -k "c3 or sample or counts or relabel or probs or shots or measure or swap_relab or golden" > gpurun_out/pytest_sampler.log 2>&1; echo "sampler tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_sampler.log | cut -c1-300 | head
python scripts/prof_sample.py 30
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/sample_launches.csv -k regex:"k_probs|k_samp" python scripts/prof_sample.py 30 > /dev/null 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open("gpurun_out/sample_launches.csv")))
h=[r for r in rows if "Kernel Name" in r][0]; i=rows.index(h); k=h.index("Kernel Name"); v=h.index("Metric Value")
for r in rows[i+1:]:
    if len(r)>v: print(r[k][:40], r[v])
PY
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "ms", j["ms_per_step"], "frac", j["roofline"]["frac"])
    e = j["e2e"]
    print("e2e", e["value"], "A", e["run_a_probs_ms"], "B", e["run_b_sample_ms"])
except Exception as ex:
    print("bench parse failed", ex)
PY
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_gpu.log | cut -c1-300 | head -20
