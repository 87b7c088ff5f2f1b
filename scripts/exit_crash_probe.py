"""Diagnosis: a short-lived process that used the background JIT pool died with SIGSEGV at exit on the GPU box.
Runs the same program under several settings and prints return code + stderr tail of each."""
import os, subprocess, sys, tempfile
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
prog = (
    "import sys, json; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
    "from catalyst_b200 import capi, workloads\nfrom dsl import apply_ops\n"
    "d = capi.B200Device(16, jit_min_qubits=10, jit_threshold=1)\n"
    "apply_ops(d, workloads.c3_qft_random(16, layers=3)); d.synchronize(); d.jit_wait()\n"
    "d.set_basis_state([0]*16, list(range(16))); apply_ops(d, workloads.c3_qft_random(16, layers=3)); st = d.state()\n"
    "s = d.stats(); d.close()\n"
    "print('ok', s.jit_compiled, s.jit_disk_hits, s.jit_passes, s.tile_passes, flush=True)\n"
) % (root, os.path.join(root, "tests"))
variants = {
    "default+handler": {"B200SV_DEBUG_SEGV": "1"},
    "sync compile (no pool)": {"B200SV_JIT": "always", "B200SV_DEBUG_SEGV": "1"},
    "pool of 1": {"B200SV_JIT_THREADS": "1", "B200SV_DEBUG_SEGV": "1"},
    "cache off": {"B200SV_JIT_CACHE": "off", "B200SV_DEBUG_SEGV": "1"},
    "jit off": {"B200SV_JIT": "off", "B200SV_DEBUG_SEGV": "1"},
    "second run (disk hits)": {"B200SV_DEBUG_SEGV": "1"},
}
cache = tempfile.mkdtemp(prefix="b200sv_probe_")
for name, extra in variants.items():
    env = dict(os.environ, B200SV_JIT_CACHE=os.path.join(cache, "c"))
    env.update(extra)
    out = subprocess.run([sys.executable, "-c", prog], env=env, capture_output=True, text=True, timeout=600)
    print(f"== {name}: rc={out.returncode} stdout={out.stdout.strip()[-80:]!r}")
    if out.returncode != 0:
        print(out.stderr[-2500:])
