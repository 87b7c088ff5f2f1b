#!/bin/bash
# round 2, GPU call J (1 GPU): pipelined D2H — parity of the paths that use it, PCIe reference number, bench
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_configs.py -m gpu -q --timeout 240 -p no:cacheprovider -k "strided or c1 or c3 or random_circuit or c5_single" > gpurun_out/pytest_d2h.log 2>&1; echo "d2h tests rc=$?"; tail -3 gpurun_out/pytest_d2h.log
python - <<'PY'
import torch, time
x = torch.empty(1 << 30, dtype=torch.float64, device="cuda")   # 8 GiB
h = torch.empty(1 << 27, dtype=torch.float64).pin_memory()      # 1 GiB pinned
torch.cuda.synchronize()
for rep in range(2):
    t0 = time.perf_counter()
    for k in range(8):
        h.copy_(x[k << 27:(k + 1) << 27], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"pinned D2H reference: 8 GiB in {dt*1e3:.1f} ms = {8.59/dt:.1f} GB/s", flush=True)
PY
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "ms", j["ms_per_step"], "frac", j["roofline"]["frac"])
    e = j["e2e"]
    print("e2e", e["value"], "A", e["run_a_probs_ms"], "B", e["run_b_sample_ms"], "split", json.dumps(e.get("host_split_ms", {}).get("run_a")), json.dumps(e.get("host_split_ms", {}).get("run_b")))
    print("extra", json.dumps(j.get("extra"))[:900])
except Exception as ex:
    print("bench parse failed", ex)
PY
