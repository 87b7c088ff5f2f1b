#!/bin/bash
# round 2, GPU call Q (1 GPU): tile-staged Pauli-word sums (Expval of Hamiltonians, phi = H psi), then C2/C4 timings
# stand-alone and inside the full bench
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_jit.py tests/test_gpu_plugin.py tests/test_gpu_zz_reference_programs.py -m gpu -q -s --timeout 240 -p no:cacheprovider -k "grad or expval or c2 or c4 or observ or golden or var or hamil or words" > gpurun_out/pytest_words.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_words.log | cut -c1-300 | head -20; grep "^\[parity\] C[24]" gpurun_out/pytest_words.log | cut -c1-200
timeout 600 python bench.py --only-extra > gpurun_out/extra_only.json 2> gpurun_out/extra_only.err; echo "only-extra rc=$?"; cat gpurun_out/extra_only.json | cut -c1-2500; tail -3 gpurun_out/extra_only.err
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
    print("value", j["value"], "e2e", j["e2e"]["value"], "frac", j["roofline"]["frac"])
    print("extra", json.dumps(j.get("extra")))
except Exception as ex:
    print("bench parse failed", ex)
PY
