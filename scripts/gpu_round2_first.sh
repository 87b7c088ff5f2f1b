#!/bin/bash
# First GPU call of round 2 (about 6 GPU-minutes on one B200):
#  1. the full GPU suite, which includes the two plugin-path tests of dynamic qubit release that were
#     written after round 1's GPU budget ended (tests/test_gpu_zz_reference_programs.py::test_runtime_*);
#  2. bench, both arms;
#  3. the register-blocking sweep on today's heavier passes (rows down to 2^3 amplitudes): 8 against 16
#     amplitudes per thread, with and without the double-buffered tile prefetch -- round 1's ncu capture
#     (profiles/r1_v7_pass_ncu_details.csv) shows the heavy QFT passes issue-bound at 12 warps per SM.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest_gpu.log
tail -3 gpurun_out/r2_pytest_gpu.log
python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/r2_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_reference_arm.json 2>> gpurun_out/r2_bench.err; echo "ref rc=$?"
timeout 300 python scripts/jit30.py 30 4:11:1:3 3:11:1:3 4:11:0:3 3:11:0:3 3:10:1:3 > gpurun_out/r2_jit30_sweep.log 2>&1; echo "sweep rc=$?"
grep "rep 3" gpurun_out/r2_jit30_sweep.log
# adjoint gradient on registers smaller than the GPU suite covers (n = 1..3): closed form -cos(theta)
python - <<'PY' > gpurun_out/r2_small_gradient.log 2>&1
import sys, numpy as np
sys.path[:0] = [".", "tests"]
from catalyst_b200 import capi
for n in (1, 2, 3):
    for th in (1.0, 2.0):
        try:
            d = capi.B200Device(n)
            d.start_tape(); d.named_op("RX", [th], [0]); d.expval(d.obs_named("PauliY", 0)); d.stop_tape()
            g = d.gradient()
            print(n, th, g, "ok" if abs(g[0, 0] + np.cos(th)) < 1e-12 else "MISMATCH")
        except Exception as e:
            print(n, th, "ERROR", e)
PY
cat gpurun_out/r2_small_gradient.log
