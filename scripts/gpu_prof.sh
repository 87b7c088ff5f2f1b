#!/bin/bash
# bench (both arms) + ncu launch list + full capture of six specialised passes (2 QFT, 4 random-layer)
mkdir -p gpurun_out
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pass -s 4 -c 6 -f -o gpurun_out/pass_full python scripts/jit30.py 30 4:11:1:3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
ls -la gpurun_out
