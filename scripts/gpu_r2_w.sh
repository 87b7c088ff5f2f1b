#!/bin/bash
# round 2, GPU call W (1 GPU): leaner k_bilinear_batch inner loop; ncu --set full of k_pauli_words_tile (both modes)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_jit.py tests/test_gpu_plugin.py tests/test_gpu_zz_reference_programs.py -m gpu -q --timeout 240 -p no:cacheprovider -k "grad or c2 or c4 or adjoint or golden" > gpurun_out/pytest_grad.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_grad.log | cut -c1-300 | head -20
timeout 600 python bench.py --only-extra > gpurun_out/extra_only.json 2> gpurun_out/extra_only.err; echo "only-extra rc=$?"; python - <<'PY'
import json
j = json.loads(open("gpurun_out/extra_only.json").read().strip().splitlines()[-1])
for k, v in j.items(): print(k, {a: (round(b, 2) if isinstance(b, float) else b) for a, b in v.items() if a.endswith("_ms")})
PY
timeout 600 ncu --set full --clock-control none -k regex:"k_pauli_words_tile" --launch-skip 138 -c 2 -f -o /tmp/words_dot python scripts/prof_grad.py c4 26 > gpurun_out/ncu_words.log 2>&1; echo "ncu dot rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_pauli_words_tile" --launch-skip 207 -c 2 -f -o /tmp/words_apply python scripts/prof_grad.py c4 26 >> gpurun_out/ncu_words.log 2>&1; echo "ncu apply rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_bilinear_batch<|k_bilinear_diag" --launch-skip 16 -c 6 -f -o /tmp/batch_full python scripts/prof_grad.py c4 26 >> gpurun_out/ncu_words.log 2>&1; echo "ncu batch rc=$?"
for f in words_dot words_apply batch_full; do ncu -i /tmp/$f.ncu-rep --page details --csv > gpurun_out/${f}_ncu_details.csv 2>/dev/null; done
ls -la gpurun_out/*_ncu_details.csv
