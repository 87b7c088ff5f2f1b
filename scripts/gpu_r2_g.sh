#!/bin/bash
# round 2, GPU call G (1 GPU): exit-crash fix, no-alloc test alone, then the whole suite with -x
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest "tests/test_gpu_jit.py::test_jit_disk_cache_across_processes" "tests/test_gpu_jit.py::test_process_exit_with_a_compilation_in_flight" -m gpu -q -p no:cacheprovider > gpurun_out/pytest_exit.log 2>&1; echo "exit tests rc=$?"
tail -25 gpurun_out/pytest_exit.log | cut -c1-400
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q -s --timeout 240 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -x rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep "passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_gpu.log | cut -c1-400 | head -30
grep -A16 "achieved relative errors" gpurun_out/pytest_gpu.log | head -20
du -sh gpurun_out
