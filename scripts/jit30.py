"""C3 at n qubits on the specialised (JIT) tile passes: `python scripts/jit30.py [n] [cfg ...]`,
cfg = regbits:tile_bits:dbuf[:min_row_bits], e.g. 4:11:1:7.  Prints per-run device time of the circuit."""
import sys, time, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
cfgs = sys.argv[2:] or ["4:11:1"]
ops = workloads.c5_random_depth(n, depth=40) if os.environ.get('JIT30_WORKLOAD') == 'c5' else workloads.c3_qft_random(n)
for cfg in cfgs:
    f = [int(x) for x in cfg.split(":")]
    rb, tb, db = f[:3]
    mr = f[3] if len(f) > 3 else 7
    d = capi.B200Device(n, jit="auto", jit_regbits=rb, tile_bits=tb, jit_dbuf=db, jit_threshold=1, min_row_bits=mr)
    for rep in range(4):
        d.set_basis_state([0] * n, list(range(n))); d.synchronize(); d.reset_stats(); d.set_timing(True)
        t0 = time.time(); apply_ops(d, ops); d.synchronize(); t1 = time.time()
        s = d.stats()
        print(cfg, "rep", rep, "passes", s.tile_passes, "jit", s.jit_passes, "rounds", s.rounds,
              "dev_ms %.1f wall_ms %.1f" % (s.gate_device_ms, (t1 - t0) * 1e3),
              "gates/s %.0f" % (s.gates_enqueued / (s.gate_device_ms * 1e-3)),
              "pass GB/s %.0f" % (s.tile_passes * 32 * 2 ** n / (s.gate_device_ms * 1e-3) / 1e9),
              "compiled", s.jit_compiled, "compile_ms %.0f" % s.jit_compile_ms, flush=True)
        t0 = time.time(); d.jit_wait(); tw = time.time() - t0
        if tw > 0.01: print("  waited %.2f s for background compiles" % tw, flush=True)
    print("expval Z0", d.expval(d.obs_named("PauliZ", 0)), flush=True)
    d.close()
