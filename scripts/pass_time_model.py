#!/usr/bin/env python
"""Per-pass time model of the specialised tile passes, fitted to the 32 passes whose
CUDA-event times were logged on a B200 at 30 qubits (`B200SV_PASS_LOG`, GPU call of round 2 with min_row_bits 1 and 2):

    ms = a * row_cost(c) + b * max(0, fp64_per_thread - 500) / 1000 + d * rounds + e

and applied to the kernels the planner + generator produce in DRY-RUN mode (no GPU): predicted circuit times for any
planner setting, single GPU or sharded (per-pass time scales with 2^(local qubits - 30)).

    python scripts/pass_time_model.py c3 30                      # planner variants on config 3
    python scripts/pass_time_model.py c5 34 8                    # config 5 at 34 qubits as rank 0 of 8 sees it

What it reproduced (measured / predicted, ms): C3 greedy 130.0 / 129.1, refinement mode 1 119.1 / 123.6, mode 2
119.8-121.3 / 121.5; C5 at 30 qubits 409.3 / 401.0 and 381.3 / 389.0; pass time of the 34-qubit runs of round 2 on
8 / 4 / 2 GPUs 996.6 / 1025, 1975.5 / 2035, 3844.2 / 3864.
"""
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (row bits c, rounds, ops, FP64 instructions per thread, measured ms)
MEASURED = [(5, 2, 160, 600, 5.907), (5, 2, 150, 1098, 7.842), (3, 2, 132, 1220, 8.466), (10, 3, 81, 1496, 10.172),
            (5, 4, 54, 1236, 10.837), (3, 3, 26, 576, 7.963), (4, 3, 24, 564, 6.128), (8, 2, 26, 548, 5.944),
            (4, 4, 50, 1116, 9.750), (2, 5, 49, 1184, 10.713), (2, 3, 32, 728, 6.768), (8, 3, 28, 656, 6.312),
            (2, 4, 38, 920, 8.693), (5, 3, 32, 576, 6.487), (8, 2, 19, 384, 6.030), (5, 2, 11, 192, 6.278),
            (9, 1, 3, 0, 5.775), (5, 2, 160, 600, 6.117), (5, 2, 150, 1098, 7.504), (8, 4, 78, 1784, 10.967),
            (6, 3, 33, 708, 6.771), (11, 3, 29, 704, 6.392), (4, 3, 35, 820, 7.218), (4, 3, 36, 784, 6.824),
            (2, 4, 42, 1064, 9.659), (3, 4, 39, 928, 7.583), (3, 3, 45, 936, 8.655), (7, 3, 25, 616, 6.060),
            (3, 3, 29, 580, 7.890), (8, 2, 13, 224, 6.043), (3, 2, 9, 128, 7.844), (10, 1, 1, 0, 5.736)]


def row_cost(c):
    return 1.0 if c >= 7 else {6: 1.12, 5: 1.2, 4: 1.3, 3: 1.38, 2: 1.6}.get(c, 3.0)


def features(c, rounds, fp64):
    return [row_cost(c), max(0, fp64 - 500) / 1000.0, rounds, 1.0]


def fit():
    A = np.array([features(c, r, f) for c, r, _, f, _ in MEASURED])
    y = np.array([m[4] for m in MEASURED])
    coef = np.linalg.lstsq(A, y, rcond=None)[0]
    return coef, float(np.sqrt(np.mean((A @ coef - y) ** 2)))


PROG = r"""
import sys
sys.path[:0] = [%r, %r]
from catalyst_b200 import capi, workloads
from dsl import apply_ops
cfg, n, world, kw = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), eval(sys.argv[4])
ops = workloads.c3_qft_random(n) if cfg == "c3" else workloads.c5_random_depth(n, depth=40)
if world > 1:
    kw.update(world_size=world, rank=0, comm="local:dry")
d = capi.B200Device(**kw)
d.allocate_qubits(n)
apply_ops(d, ops)
d.synchronize()
d.close()
""" % (ROOT, os.path.join(ROOT, "tests"))


def predict(cfg, n, world, kw, coef):
    env = dict(os.environ, B200SV_DRYRUN="1", B200SV_JIT_CACHE="off", B200SV_JIT_VERBOSE="1",
               B200SV_JIT_DUMP="/tmp/b200sv_model_dump")
    out = subprocess.run([sys.executable, "-c", PROG, cfg, str(n), str(world), repr(kw)], capture_output=True, text=True,
                         env=env, timeout=3600)
    P = [tuple(int(x) for x in m.groups())
         for m in re.finditer(r"c=(\d+) rounds=(\d+) ops=(\d+) fp64/thread=(\d+)", out.stdout + out.stderr)]
    if not P:
        raise RuntimeError(out.stderr[-2000:])
    nl = n - (world.bit_length() - 1)
    X = np.array([features(c, r, f) for c, r, _, f in P])
    return len(P), float((X @ coef).sum()) * 2.0 ** (nl - 30)


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "c3"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    world = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    coef, rms = fit()
    print("fit: ms = %.2f * row_cost + %.2f * max(0, fp64 - 500) / 1000 + %.2f * rounds + %.2f   (rms %.2f ms per pass)"
          % (*coef, rms))
    for label, kw in (("rows >= 128 B, greedy (round-2 bench runs)", dict(min_row_bits=3, plan_refine=0)),
                      ("64-byte rows, greedy", dict(plan_refine=0)),
                      ("refinement mode 1", dict(plan_refine_mode=1)),
                      ("refinement mode 2 (HEAD)", dict())):
        passes, ms = predict(cfg, n, world, kw, coef)
        print(f"{cfg} n={n} world={world}  {label:44s} {passes:3d} passes  predicted {ms:8.1f} ms of passes", flush=True)


if __name__ == "__main__":
    main()
