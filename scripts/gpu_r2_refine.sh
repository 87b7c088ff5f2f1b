#!/bin/bash
# round 2, last GPU call (1 GPU): planner refinement -- C3 / C5 at 30 qubits with and without, then parity with it forced on
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for pr in 0 6; do
  echo "== B200SV_PLAN_REFINE=$pr"
  B200SV_PLAN_REFINE=$pr timeout 100 python scripts/jit30.py 30 4:11:1:2 2>&1 | grep "rep 3" | cut -c1-130
  B200SV_PLAN_REFINE=$pr JIT30_WORKLOAD=c5 timeout 100 python scripts/jit30.py 30 4:11:1:2 2>&1 | grep "rep 3" | cut -c1-130
done
B200SV_PLAN_REFINE=6 timeout 130 python -m pytest tests/test_gpu_configs.py tests/test_gpu_jit.py tests/test_gpu_sharded.py -m gpu -q -x --timeout 100 -p no:cacheprovider -k "c3 or c5 or c1 or c2_live or jit_matches or random or state_random or gradient" > gpurun_out/pytest_refine.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_refine.log | cut -c1-300 | head
