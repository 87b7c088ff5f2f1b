"""One pass of config C3 (30-qubit QFT + random layers) for ncu: `python scripts/prof_c3.py [n] [reps]`."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
kw = {}
if len(sys.argv) > 3: kw["kernel_gen"] = int(sys.argv[3])
if len(sys.argv) > 4: kw["tile_bits"] = int(sys.argv[4])
d = capi.B200Device(n, **kw)
ops = workloads.c3_qft_random(n)
for rep in range(reps):
    d.set_basis_state([0] * n, list(range(n))); d.synchronize(); d.reset_stats(); d.set_timing(True)
    apply_ops(d, ops); d.synchronize()
    s = d.stats()
    print("gates", s.gates_enqueued, "passes", s.tile_passes, "rounds", s.rounds, "ms %.2f" % s.gate_device_ms, flush=True)
print("expval Z0", d.expval(d.obs_named("PauliZ", 0)))
