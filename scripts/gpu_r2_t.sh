#!/bin/bash
# round 2, GPU call T (1 GPU): distributed sampler + slice-mode State/Probs on virtual shards, then the sharded suite
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q -x --timeout 240 -p no:cacheprovider -k "distributed or slices" > gpurun_out/pytest_dist.log 2>&1; echo "dist tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_dist.log | cut -c1-400 | head -30
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_configs.py -m gpu -q --timeout 240 -p no:cacheprovider > gpurun_out/pytest_sharded.log 2>&1; echo "sharded suite rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_sharded.log | cut -c1-300 | head -20
