#!/bin/bash
# compute-sanitizer over smoke() and a 20-test subset with every pass on run-time specialised kernels
# (B200SV_JIT=always): memcheck (out-of-bounds / misaligned accesses, leaks at exit) and racecheck (shared-memory
# hazards between the rounds of a tile pass).  Logs -> gpurun_out/sanitizer_*.log
mkdir -p gpurun_out
export B200SV_JIT_CACHE=off PYTHONUNBUFFERED=1
SUBSET="tests/test_gpu_jit.py::test_jit_random_circuit_state tests/test_gpu_jit.py::test_jit_uniform_diagonal_factors"
K='(cfg0 or 0-14-8-1 or 1-14-8-1) and (3-1 or 6-2 or 9-3 or 12-4 or 14-8-1)'
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
    python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanitizer_${tool}_smoke.log 2>&1
  echo "$tool smoke rc=$?" | tee -a gpurun_out/sanitizer_${tool}_smoke.log
  B200SV_JIT=always timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
    python -m pytest $SUBSET tests/test_gpu_parity.py -m gpu -q -x -k "$K or golden or measure" \
    > gpurun_out/sanitizer_${tool}_tests.log 2>&1
  echo "$tool tests rc=$?" | tee -a gpurun_out/sanitizer_${tool}_tests.log
  grep -E "ERROR SUMMARY|passed|failed" gpurun_out/sanitizer_${tool}_*.log
done
