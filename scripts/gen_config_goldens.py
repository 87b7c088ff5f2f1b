#!/usr/bin/env python
"""Generate tests/golden/configs/*.npz: results of BASELINE.json's configs C2-C5 computed by the CPU
oracle (oracle/, the pinned restatement of the reference algorithm) at register sizes where the oracle
needs minutes, so that the GPU parity tests (tests/test_gpu_configs.py) can check the CUDA device at
those sizes without spending that time on the GPU box.  Only small digests travel: expectation values,
Jacobians, probes of the state / probability vectors at fixed pseudo-random indices, the SHA-256 of the
sampled basis states.

Run in the build container (CPU only):  python scripts/gen_config_goldens.py [c2 c3 c4 c5]
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import config_cases as cc  # noqa: E402
from oracle import oracle as om  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "configs")

# (case, register size, extra kwargs): the sizes the fixtures are made at
CASES = {
    "c2": (cc.run_c2, 24, {}),
    "c3": (cc.run_c3, 26, {}),
    "c4": (cc.run_c4, 22, {}),
    "c5": (cc.run_c5, 26, {}),
}


def main():
    om.build()
    os.makedirs(OUT, exist_ok=True)
    which = sys.argv[1:] or list(CASES)
    for name in which:
        fn, n, kw = CASES[name]
        t0 = time.time()
        dev = om.OracleDevice(n, seed=cc.SEED)
        res = fn(dev, n, **kw)
        del dev
        keep = {k: np.asarray(res[k]) for k in cc.FIXTURE_KEYS[name]}
        keep["n"] = np.int64(n)
        path = os.path.join(OUT, f"{name}_n{n}.npz")
        np.savez_compressed(path, **keep)
        print(f"{name}: n={n} -> {path} ({os.path.getsize(path)} bytes, {time.time() - t0:.0f} s, "
              f"{om.num_threads()} threads)", flush=True)


if __name__ == "__main__":
    main()
