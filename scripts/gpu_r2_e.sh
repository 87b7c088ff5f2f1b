#!/bin/bash
# round 2, GPU call E (1 GPU): the no-alloc entry point alone in a fresh process, then the whole suite with -x
# (the no-alloc file is collected last), then bench
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
tail -15 gpurun_out/pytest_noalloc_alone.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 240 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  \|native backtrace" gpurun_out/pytest_gpu.log | cut -c1-400 | head -30
grep -A12 "native backtrace" gpurun_out/pytest_gpu.log | cut -c1-300 | head -40
grep "^\[parity\]" gpurun_out/pytest_gpu.log | cut -c1-220 | head -40
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    for k in ("value", "ms_per_step", "e2e", "cold_start", "extra"):
        print(k, json.dumps(j.get(k))[:1200])
except Exception as e:
    print("bench parse failed", e)
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
du -sh gpurun_out
