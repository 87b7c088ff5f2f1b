#!/bin/bash
# round 2, GPU call H (8 GPUs): NCCL parity at 2/4/8 ranks (push / staged / in-place swaps), bench at N=8 with the
# push-mode and the staged swap, N=4, N=2, reference arm at N=8
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader | sort | uniq -c > gpurun_out/gpu8.txt
timeout 900 python -m pytest tests/test_gpu_nccl.py -m gpu -q -s --timeout 300 -p no:cacheprovider > gpurun_out/pytest_nccl_n8.log 2>&1; echo "nccl pytest rc=$?" | tee -a gpurun_out/pytest_nccl_n8.log
grep "nccl parity\|passed\|failed\|FAILED\|Timeout\|^E  " gpurun_out/pytest_nccl_n8.log | sort -u | cut -c1-250 | head -40
run_bench() { # N push tag
  B200SV_SWAP_PUSH=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 2960$1 \
    bench.py --gpus $1 --steps 2 --warmup 3 > gpurun_out/bench_n$1_$3.json 2> gpurun_out/bench_n$1_$3.err; echo "bench n$1 $3 rc=$?"
  python - <<PY
import json
try:
    j = json.loads([l for l in open("gpurun_out/bench_n$1_$3.json") if l.startswith("{")][-1])
    print("  circuit_seconds", j.get("circuit_seconds"), "value", j.get("value"), "e2e", j["e2e"]["value"])
    print("  sharding", json.dumps(j.get("sharding"))[:700])
    print("  parity_check", json.dumps(j.get("parity_check"))[:400])
    print("  roofline frac", j["roofline"]["frac"], "bound", j["roofline"]["bound"], "fp64_frac", j["roofline"]["fp64_frac"], "avg_ms", j["roofline"]["avg_launch_ms"])
    print("  cold", j["cold_start"]["first_step_ms"], "nvrtc_cpu_s", j["cold_start"]["nvrtc_cpu_s"], "disk", j["cold_start"]["first_step_kernels_from_disk_cache"])
except Exception as e:
    print("  parse failed", e)
PY
  tail -2 gpurun_out/bench_n$1_$3.err | cut -c1-300
}
run_bench 8 1 push
run_bench 8 0 staged
run_bench 4 1 push
run_bench 2 1 push
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/bench_ref_n8.json 2> gpurun_out/bench_ref_n8.err; echo "ref n8 rc=$?"
grep "^{" gpurun_out/bench_ref_n8.json | cut -c1-500
du -sh gpurun_out
