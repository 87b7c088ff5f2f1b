"""Where does the time of the terminal measurements go on a sharded register?  Virtual shards on one GPU:
`python scripts/prof_sharded_meas.py [n] [world]`."""
import sys, time, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 28
world = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ops = workloads.c5_random_depth(n, depth=6)


def prog(d, r):
    d.allocate_qubits(n)
    out = []
    for rep in range(3):
        d.set_basis_state([0] * n, list(range(n)))
        t0 = time.perf_counter(); apply_ops(d, ops); d.synchronize(); t1 = time.perf_counter()
        ez = d.expval(d.obs_named("PauliZ", 0)); t2 = time.perf_counter()
        pp = d.probs(list(range(10))); t3 = time.perf_counter()
        pp2 = d.probs(list(range(10))); t4 = time.perf_counter()
        s = d.stats()
        out.append("rank %d rep %d circuit %.1f ms expval %.1f ms probs %.1f ms probs-again %.1f ms exchanges %d layout %s" % (
            r, rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), s.exchanges, d.layout()))
    return out


for lines in capi.run_spmd(world, prog):
    print("\n".join(lines[-1:]))
