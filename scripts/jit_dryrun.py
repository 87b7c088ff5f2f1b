#!/usr/bin/env python
"""Development tool (no GPU needed): plan a workload into tile passes, generate the specialised kernels and
compile them with NVRTC for sm_100a — B200SV_DRYRUN=1 makes the device library skip every CUDA call.
Sources and cubins land in $B200SV_JIT_DUMP (default gpurun_out/jit_dump); the per-kernel register count
and SASS instruction mix are then read with cuobjdump (see scripts/sass_mix.py).

    python scripts/jit_dryrun.py c3 30        # config, qubits
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
os.environ["B200SV_DRYRUN"] = "1"
os.environ.setdefault("B200SV_JIT_DUMP", os.path.join(ROOT, "gpurun_out", "jit_dump"))
os.environ.setdefault("B200SV_JIT_CACHE", "off")
os.environ.setdefault("B200SV_JIT_VERBOSE", "1")

from catalyst_b200 import capi, workloads  # noqa: E402
from dsl import apply_ops  # noqa: E402


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "c3"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    if cfg == "c3":
        ops = workloads.c3_qft_random(n)
    elif cfg == "c5":
        ops = workloads.c5_random_depth(n, depth=int(sys.argv[3]) if len(sys.argv) > 3 else 40)
    elif cfg == "c2":
        ops = workloads.c2_qaoa_maxcut(n)[0]
    elif cfg == "c4":
        ops = workloads.c4_vqe(n)[0]
    else:
        raise SystemExit("unknown config")
    d = capi.B200Device(n)
    apply_ops(d, ops)
    d.flush()
    s = d.stats()
    amps = float(1 << n)
    print(f"{cfg} n={n}: {s.gates_enqueued} gates, {s.tile_passes} passes ({s.jit_passes} specialised), "
          f"{s.rounds} rounds, {s.kernel_ops} canonical ops, fp64 instr per amplitude per pass "
          f"{s.fp64_ops / amps / max(1, s.jit_passes):.1f}")


if __name__ == "__main__":
    main()
