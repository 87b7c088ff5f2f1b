#!/bin/bash
# round 2, last GPU call (1 GPU): planner refinement forced on -- C3 / C5 at 30 qubits, then parity
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
export B200SV_PLAN_REFINE=6
timeout 40 python scripts/jit30.py 30 4:11:1:2 2>&1 | grep "rep 3" | cut -c1-130
JIT30_WORKLOAD=c5 timeout 60 python scripts/jit30.py 30 4:11:1:2 2>&1 | grep "rep 3" | cut -c1-130
timeout 60 python -m pytest tests/test_gpu_configs.py tests/test_gpu_jit.py -m gpu -q -x --timeout 50 -p no:cacheprovider -k "c3_live or c5_single or c1 or c2_live or jit_matches or c5_virtual_shards_live" > gpurun_out/pytest_refine.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_refine.log | cut -c1-300 | head
