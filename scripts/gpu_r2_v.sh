#!/bin/bash
# round 2, GPU call V (1 GPU): full suite (-x, no-alloc test last), bench, and one `ncu --set full` capture of the
# gradient / Hamiltonian kernels (C4, 26 qubits; launches of the second repetition)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_gpu.log | cut -c1-300 | head -20
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "ms", j["ms_per_step"], "frac", j["roofline"]["frac"], "launches", j["gpu_launches"])
    e = j["e2e"]
    print("e2e", e["value"], {k: v for k, v in e.items() if k.endswith("_ms")})
    print("cpu", j["cpu_baseline"]["value"] if j.get("cpu_baseline") else None, "cold", json.dumps(j.get("cold_start"))[:300])
    print("extra", json.dumps(j.get("extra"))[:1500])
except Exception as ex:
    print("bench parse failed", ex)
PY
timeout 600 ncu --set full --clock-control none -k regex:"k_pauli_words_tile|k_bilinear_batch|k_bilinear_diag_batch" --launch-skip 150 -c 8 -f -o /tmp/grad_full python scripts/prof_grad.py c4 26 > gpurun_out/ncu_grad_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/grad_full.ncu-rep --page details --csv > gpurun_out/grad_ncu_details.csv 2>/dev/null
ncu -i /tmp/grad_full.ncu-rep --page raw --csv > gpurun_out/grad_ncu_raw.csv 2>/dev/null
ls -la gpurun_out/grad_ncu_*.csv; du -sh gpurun_out
