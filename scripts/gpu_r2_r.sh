#!/bin/bash
# round 2, GPU call R+S (1 GPU): double-buffered k_pauli_words_tile and k_bilinear_batch, streaming diagonal-generator batches; launch lists of
# the C4 / C2 gradients
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_jit.py tests/test_gpu_plugin.py tests/test_gpu_zz_reference_programs.py -m gpu -q -s --timeout 240 -p no:cacheprovider -k "grad or expval or c2 or c4 or observ or golden or var or hamil or words or adjoint" > gpurun_out/pytest_words.log 2>&1; echo "tests rc=$?"; grep "passed\|failed\|FAILED\|^E  " gpurun_out/pytest_words.log | cut -c1-300 | head -20; grep "^\[parity\] C[24]" gpurun_out/pytest_words.log | cut -c1-200
timeout 600 python bench.py --only-extra > gpurun_out/extra_only.json 2> gpurun_out/extra_only.err; echo "only-extra rc=$?"; cat gpurun_out/extra_only.json | cut -c1-2500; tail -3 gpurun_out/extra_only.err
for cfg in "c4 26" "c2 28"; do
  set -- $cfg
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/grad_launches_$1.csv python scripts/prof_grad.py $1 $2 > /dev/null 2>&1
  python - gpurun_out/grad_launches_$1.csv <<'PY'
import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1])))
h=[r for r in rows if "Kernel Name" in r][0]; i=rows.index(h); k=h.index("Kernel Name"); v=h.index("Metric Value")
tot=collections.Counter(); cnt=collections.Counter()
data=[(r[k].split("(")[0][:40], float(r[v].replace(",",""))) for r in rows[i+1:] if len(r)>v]
half=len(data)//2
for name,t in data[half:]:
    name = "k_pass_*" if name.startswith("k_pass_") else name
    tot[name]+=t; cnt[name]+=1
print(sys.argv[1], "second half of the launch list (rep 1), total %.1f ms" % (sum(tot.values())/1e6))
for name,t in tot.most_common(12): print("  %-42s n=%4d total %8.2f ms avg %.3f ms"%(name,cnt[name],t/1e6,t/1e6/cnt[name]))
PY
done
