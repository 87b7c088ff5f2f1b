#!/bin/bash
# round 2, GPU call A (1 GPU): parity suite incl. the config-size tests, variant sweep of the pass generator,
# bench (both arms), launch list + full ncu capture of the specialised passes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt
nproc >> gpurun_out/gpu.txt; free -g | head -2 >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -s -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -c "^\[parity\]" gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log
export B200SV_JIT_CACHE=off
for v in "1 4:11:1:3" "0 4:11:1:3" "1 3:11:1:3" "1 4:11:1:7"; do
  set -- $v
  echo "== B200SV_JIT_UNIFORM=$1 cfg=$2" >> gpurun_out/sweep.log
  B200SV_JIT_UNIFORM=$1 timeout 600 python scripts/jit30.py 30 $2 >> gpurun_out/sweep.log 2>&1
done
tail -40 gpurun_out/sweep.log
unset B200SV_JIT_CACHE
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pass -c 18 -f -o gpurun_out/pass_full python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -20
