"""C4 (or C2) tape + expval + adjoint gradient, twice (second run on compiled kernels); under
`ncu --metrics gpu__time_duration.sum` gives the launch list of the gradient's kernels."""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops, build_obs
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else (26 if cfg == "c4" else 28)
ops, obs = workloads.c4_vqe(n, 4, 1000, seed=0) if cfg == "c4" else workloads.c2_qaoa_maxcut(n, 6, seed=0)
d = capi.B200Device(n, seed=37)
for rep in range(2):
    d.set_basis_state([0] * n, list(range(n))); d.synchronize()
    d.start_tape(); apply_ops(d, ops); d.synchronize()
    t0 = time.perf_counter(); ev = d.expval(build_obs(d, obs)); t1 = time.perf_counter()
    d.stop_tape()
    d.reset_stats()
    g = d.gradient(); t2 = time.perf_counter()
    s = d.stats()
    print(cfg, n, "rep", rep, "expval ms %.1f gradient ms %.1f" % (1e3 * (t1 - t0), 1e3 * (t2 - t1)), "passes", s.tile_passes, "jit", s.jit_passes,
          "other launches", s.other_launches, flush=True)
    d.jit_wait()
d.close()
