"""C3 at n qubits, then Probs-on-device + Sample(1e5): run under `ncu --metrics gpu__time_duration.sum` for the
launch list of the measurement kernels (k_probs_all*, k_samp_*)."""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi, workloads
from dsl import apply_ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
d = capi.B200Device(n, seed=37)
ops = workloads.c3_qft_random(n)
for rep in range(2):
    d.set_basis_state([0] * n, list(range(n)))
    apply_ops(d, ops); d.synchronize(); d.jit_wait()
    t0 = time.perf_counter(); s = d.sample(100000); t1 = time.perf_counter()
    print("rep", rep, "sample ms %.2f mean %.5f" % (1e3 * (t1 - t0), s.mean()), flush=True)
d.close()
