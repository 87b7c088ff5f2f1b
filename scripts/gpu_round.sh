#!/bin/bash
# one GPU call: parity tests, bench (both arms), launch list, one full ncu capture of the tile kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/jit30.py 30 4:11:1 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
# launches 0-33 are the interpreted first run (2 x k_set_one + 32 x k_tile3); the specialised kernels follow
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pass -c 34 -f -o gpurun_out/pass_full python scripts/jit30.py 30 4:11:1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
