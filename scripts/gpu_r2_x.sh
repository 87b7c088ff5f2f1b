#!/bin/bash
# round 2, GPU call X (1 GPU): gradient work vectors kept between calls; C2/C4 timings three times over
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_plugin.py tests/test_gpu_zz_reference_programs.py -m gpu -q --timeout 240 -p no:cacheprovider -k "grad or c2 or c4 or adjoint or golden or keep or pressure or release" > gpurun_out/pytest_grad.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_grad.log | cut -c1-300 | head -20
for i in 1 2 3; do
timeout 600 python bench.py --only-extra > gpurun_out/extra_only_$i.json 2> gpurun_out/extra_only.err; python - gpurun_out/extra_only_$i.json <<'PY'
import json, sys
j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k: {a: round(b, 1) for a, b in v.items() if a.endswith("_ms") and not a.startswith("per_")} for k, v in j.items()})
PY
done
