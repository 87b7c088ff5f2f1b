import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
for cfg in [dict(), dict(fuse=False), dict(tile_bits=12), dict(tile_bits=10)]:
    d = capi.B200Device(n, **cfg)
    ops = workloads.c3_qft_random(n)
    for name, sub in [("qft", workloads.qft(n)), ("layers", workloads.random_layers(n, 10, 0)), ("c3", ops)]:
        if cfg.get("fuse") is False and name != "layers": continue
        d.reset(n); d.synchronize(); d.reset_stats(); d.set_timing(True)
        t0 = time.time(); apply_ops(d, sub); d.synchronize(); t1 = time.time()
        s = d.stats()
        print(cfg, name, "gates", s.gates_enqueued, "passes", s.tile_passes, "ops", s.kernel_ops, "rounds", s.rounds, "dev_ms %.1f wall_ms %.1f" % (s.gate_device_ms, (t1-t0)*1e3),
              "gates/s %.0f" % (s.gates_enqueued / (s.gate_device_ms*1e-3)), "alg GB/s %.0f" % (s.algorithmic_bytes/ (s.gate_device_ms*1e-3)/1e9),
              "pass GB/s %.0f" % (s.tile_passes * 32 * 2**n / (s.gate_device_ms*1e-3)/1e9), flush=True)
    t0=time.time(); p = d.expval(d.obs_named("PauliZ", 0)); print("expval", p, time.time()-t0)
    t0=time.time(); smp = d.sample(100000); print("sample", smp.shape, smp.mean(), time.time()-t0)
    d.close()
