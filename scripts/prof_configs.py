"""Wall-clock of BASELINE.json configs 1, 2 and 4 on the device (config 3 / 5 are bench.py's):
`python scripts/prof_configs.py [c1 c2 c4 ...] [--reps R]`.  Each config runs R times (default 3: the
first interpreted, later ones on specialised passes); prints one line per repetition."""
import sys, time, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
from catalyst_b200 import capi, workloads
from dsl import apply_ops, build_obs
args = [a for a in sys.argv[1:] if not a.startswith("--")]
reps = int(os.environ.get("REPS", 3))
which = args or ["c1", "c2", "c4"]
n2 = int(os.environ.get("C2_QUBITS", 28)); n4 = int(os.environ.get("C4_QUBITS", 26))


def run(name, n, ops, obs, grad, mcm=False):
    kw = eval(os.environ.get("B200_KW", "{}"))
    d = capi.B200Device(n, seed=37, **kw)
    for rep in range(reps):
        d.set_basis_state([0] * n, list(range(n))); d.synchronize(); d.reset_stats()
        t0 = time.perf_counter()
        if grad: d.start_tape()
        if mcm:
            half = len(ops) // 2
            apply_ops(d, ops[:half])
            m = [d.measure(w) for w in range(4)]
            apply_ops(d, ops[half:])
        else:
            apply_ops(d, ops)
        d.synchronize(); t1 = time.perf_counter()
        e = d.expval(build_obs(d, obs)); t2 = time.perf_counter()
        s = d.stats()
        g = None
        if grad:
            d.stop_tape()
            g = d.gradient(); t3 = time.perf_counter()
        line = f"{name} n={n} rep {rep}: circuit {1e3*(t1-t0):.1f} ms ({len(ops)} gates, {s.tile_passes} passes, {s.jit_passes} specialised)  expval {1e3*(t2-t1):.1f} ms = {e:.12f}"
        if grad: line += f"  gradient {1e3*(t3-t2):.1f} ms ({g.size} entries, |g|={np.linalg.norm(g):.6f}, launches {d.stats().other_launches - s.other_launches}, passes {d.stats().tile_passes - s.tile_passes})"
        if mcm: line += f"  mcm {m}"
        print(line, flush=True)
        d.jit_wait()
    d.close()


if "c1" in which:
    ops, obs = workloads.c1_strongly_entangling()
    run("C1", 20, ops, obs, False)
if "c2" in which:
    ops, obs = workloads.c2_qaoa_maxcut(n=n2)
    run("C2", n2, ops, obs, True)
if "c4" in which:
    ops, obs = workloads.c4_vqe(n=n4)
    run("C4-expval+grad", n4, ops, obs, True)
    run("C4-mcm", n4, ops, ("N", "PauliZ", 5), False, mcm=True)
