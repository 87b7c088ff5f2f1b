#!/usr/bin/env python
"""bench.py — BASELINE.json metric: gates/s of the 30-qubit circuit on one B200 (+ HBM fraction).

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W [--impl reference]`
prints ONE JSON line on rank 0.

  step      = one execution of the workload's whole gate sequence from |0...0> on a state resident
              in HBM (config 3 of BASELINE.json: 30-qubit QFT + 10 random 1q/2q layers, 930 gates)
  value     = gates/s, whole job, timed on the device with the state already in HBM
  e2e       = same metric through the C ABI with HOST buffers: host gate parameters in, 1e5 shots
              of samples out (the sampler's uniforms go host->device, the indices device->host)
  roofline  = the fused tile-pass kernel: algorithmic bytes (2*16*2^(n-c) per gate)
              / CUDA-event time of the passes, against the measured HBM copy bandwidth
  cpu_baseline / --impl reference = the CPU oracle (a port of the reference algorithm; the real
              lightning.qubit is an absent third-party package) on all host cores, bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

SHOTS = 100_000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--qubits", type=int, default=0, help="override the register size (default: 30 at N=1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-gates", type=int, default=16)
    ap.add_argument("--tile-bits", type=int, default=0)
    return ap.parse_args()


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def stride_sample(ops, m):
    """A bounded, deterministic sample of the workload: m gates spread uniformly over the circuit."""
    if m >= len(ops):
        return list(ops)
    return [ops[(i * len(ops)) // m] for i in range(m)]


def cpu_reference_step(o, sample_ops):
    from dsl import apply_ops
    t0 = time.perf_counter()
    apply_ops(o, sample_ops)
    return time.perf_counter() - t0


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm's CPU implementation on the host cores."""
    if rank != 0:
        return
    from catalyst_b200 import workloads
    from oracle import oracle as om
    om.build()
    n = args.qubits or 30
    ops = workloads.c3_qft_random(n)
    sample = stride_sample(ops, args.cpu_sample_gates)
    cores = om.num_threads()
    o = om.OracleDevice(n)
    for _ in range(args.warmup):
        cpu_reference_step(o, sample[:2])
    t = 0.0
    for _ in range(args.steps):
        t += cpu_reference_step(o, sample)
    value = len(sample) * args.steps / t
    desc = (f"{len(sample)} of {len(ops)} gates (uniform stride over the circuit) per step on a "
            f"{n}-qubit state, OpenMP {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": "gates_per_second", "value": value, "unit": "gates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 (complex128)", "data": "synthetic",
        "config": {"workload": f"C3: {n}-qubit QFT + 10 random 1q/2q layers ({len(ops)} gates)",
                   "qubits": n},
        "cpu_baseline": {"value": value, "unit": "gates/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": "gates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "lightning.qubit (pennylane-lightning 0.46.0-dev27) is not installable here; this is "
                "the repo's C++/OpenMP restatement of its algorithm (oracle/), kind=port"}), flush=True)


def run_ours(args, rank, world, local_rank):
    import numpy as np
    from catalyst_b200 import capi, workloads
    from dsl import apply_ops

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n = args.qubits or 30
    ops = workloads.c3_qft_random(n)
    ngates = workloads.gate_count(ops)
    kw = {"device": local_rank}
    if args.tile_bits:
        kw["tile_bits"] = args.tile_bits
    d = capi.B200Device(n, seed=37, **kw)
    zeros = [0] * n
    wires = list(range(n))

    def step():
        d.set_basis_state(zeros, wires)  # back to |0...0> in HBM (memset, part of the step)
        apply_ops(d, ops)
        d.flush()

    def barrier():
        d.synchronize()
        if dist is not None:
            dist.barrier()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    d.reset_stats()
    d.set_timing(True)
    clocks = ClockSampler(local_rank)
    clocks.start()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    barrier()
    t1 = time.perf_counter()
    clk = clocks.stop()
    s = d.stats()
    d.set_timing(False)
    elapsed = t1 - t0
    if dist is not None:
        import torch
        tt = torch.tensor([elapsed], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        elapsed = float(tt.item())
    value = world * ngates * args.steps / elapsed
    dev_s = s.gate_device_ms * 1e-3
    peak, peak_src = measured_peak()
    achieved = s.algorithmic_bytes / dev_s / 1e9
    pass_gbs = s.tile_passes * 32.0 * (1 << n) / dev_s / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "tile_kernel_traffic.json")))["bytes_per_launch"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_tile_program", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "launches": int(s.tile_passes), "avg_launch_ms": s.gate_device_ms / max(1, s.tile_passes),
                "algorithmic_bytes_per_launch": s.algorithmic_bytes / max(1, s.tile_passes),
                "hbm_bytes_per_launch_min": 32.0 * (1 << n), "pass_gbs": pass_gbs, "pass_frac": pass_gbs / peak,
                "gates_per_launch": s.gates_enqueued / max(1, s.tile_passes)}
    launches = int(s.tile_passes + s.other_launches)

    # ---- e2e through the C ABI with host buffers: gates in, 1e5 shots out ---------------------
    e2e = None
    if rank == 0 or world > 1:
        d.reset_stats()
        t0 = time.perf_counter()
        h2d = d2h = 0
        for _ in range(args.steps):
            d.set_basis_state(zeros, wires)
            apply_ops(d, ops)
            smp = d.sample(SHOTS)  # host numpy array (shots x n doubles) written by the device lib
            st = d.stats()
        te = time.perf_counter() - t0
        st = d.stats()
        h2d = (16 * SHOTS * args.steps + st.tile_passes * 3880) // args.steps
        d2h = 8 * SHOTS
        e2e = {"value": world * ngates * args.steps / te, "unit": "gates/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * te / args.steps,
               "includes": f"|0> reset, {ngates} gate calls with host parameters, Sample({SHOTS} shots) into a host array",
               "sample_mean": float(smp.mean())}
    d.close()

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import oracle as om
        om.build()
        sample = stride_sample(ops, args.cpu_sample_gates)
        o = om.OracleDevice(n)
        cpu_reference_step(o, sample[:2])
        tc = cpu_reference_step(o, sample)
        cpu = {"value": len(sample) / tc, "unit": "gates/s", "cores": om.num_threads(), "kind": "port",
               "sample": f"{len(sample)} of {len(ops)} gates (uniform stride) on a {n}-qubit state, {tc:.1f} s"}
        del o
    if rank == 0:
        print(json.dumps({
            "metric": "gates_per_second", "value": value, "unit": "gates/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * elapsed / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 (complex128)", "data": "synthetic",
            "config": {"workload": f"C3: {n}-qubit QFT + 10 random 1q/2q layers ({ngates} gates), state resident in HBM",
                       "qubits": n, "state_bytes": 16 << n, "l2_policy": "state (>= 16 GiB at 30q) is far larger than L2",
                       "parallelism": "single GPU" if world == 1 else f"{world} independent replicas"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches,
            "clocks": clk}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
