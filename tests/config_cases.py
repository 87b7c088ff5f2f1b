"""BASELINE.json's five configs as device-agnostic test programs.

Each `run_cN(dev, n, ...)` drives an object with the QuantumDevice method set (oracle.OracleDevice or
catalyst_b200.capi.B200Device) through the exact `catalyst_b200.workloads.cN` generator and returns a
dict of numpy results.  The same function produces the golden fixtures (scripts/gen_config_goldens.py,
CPU oracle, run in the build container) and the device side of tests/test_gpu_configs.py, so both arms
run the same calls in the same order (which matters for the RNG stream: Sample and Measure draw from the
device generator seeded with SEED).

Reference programs these mirror: frontend/test/pytest/test_template.py:156,620 (templates + expval),
test_gradient.py:608-676 (adjoint Jacobians), test_mid_circuit_measurement.py:88-188 (measure + reset),
test_measurements_results.py (probs / sample layout).
"""
from __future__ import annotations

import hashlib

import numpy as np

from catalyst_b200 import workloads
from dsl import apply_ops, build_obs

SEED = 37
SHOTS = 100_000
N_PROBE = 4096


def probe_indices(n, count=N_PROBE, seed=1234):
    """Deterministic sample of basis-state indices at which large vectors are compared with a fixture."""
    rng = np.random.default_rng(seed + n)
    idx = np.unique(np.concatenate([rng.integers(0, 1 << n, size=count), [0, (1 << n) - 1]]))
    return idx.astype(np.int64)


def samples_to_indices(smp):
    """(shots, k) array of 0.0/1.0 (first wire first) -> basis-state index per shot (first wire = MSB)."""
    k = smp.shape[1]
    w = (1 << np.arange(k - 1, -1, -1)).astype(np.int64)
    return (smp.astype(np.int64) * w).sum(axis=1)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_c1(dev, n=20, layers=4):
    ops, obs = workloads.c1_strongly_entangling(n=n, layers=layers, seed=0)
    apply_ops(dev, ops)
    out = {"expval": np.float64(dev.expval(build_obs(dev, obs)))}
    st = dev.state()
    out["state_probe"] = st[probe_indices(n)]
    out["state"] = st
    return out


def run_c2(dev, n, layers=6, state=True):
    ops, ham = workloads.c2_qaoa_maxcut(n=n, layers=layers, seed=0)
    dev.start_tape()
    apply_ops(dev, ops)
    out = {"expval": np.float64(dev.expval(build_obs(dev, ham)))}
    dev.stop_tape()
    out["grad"] = dev.gradient()
    if state:
        st = dev.state()
        out["state_probe"] = st[probe_indices(n)]
        out["state"] = st
    return out


def run_c3(dev, n, layers=10, repeats=1):
    """QFT + random layers, then run A: Probs() of all wires, run B: SHOTS samples (seeded stream).
    `repeats` > 1 re-runs the circuit from |0..0> first (device side: the pass structures recur, so
    jit='auto' has specialised kernels by the last run)."""
    ops = workloads.c3_qft_random(n, layers=layers, seed=0)
    for _ in range(repeats - 1):
        apply_ops(dev, ops)
        dev.probs([0])  # flush + observe
        if hasattr(dev, "jit_wait"):
            dev.jit_wait()
        dev.set_basis_state([0] * n, list(range(n)))
    apply_ops(dev, ops)
    p = dev.probs()
    idx = samples_to_indices(dev.sample(SHOTS))
    return {"probs": p, "probs_probe": p[probe_indices(n)], "probs_sum": np.float64(p.sum()),
            "sample_idx": idx, "sample_sha256": digest(idx), "sample_head": idx[:512].copy()}


def run_c4(dev, n, layers=4, terms=1000, grad=True):
    """VQE ansatz: run A expval of the `terms`-term Hamiltonian, run B its adjoint Jacobian, run C the
    mid-circuit-measurement leg (Measure wires 0..3 on the seeded stream, then expectation values)."""
    ops, ham = workloads.c4_vqe(n=n, layers=layers, terms=terms, seed=0)
    dev.start_tape()
    apply_ops(dev, ops)
    out = {"expval": np.float64(dev.expval(build_obs(dev, ham)))}
    dev.stop_tape()
    if grad:
        out["grad"] = dev.gradient()
    # run C (same state): four mid-circuit measurements, a conditional X (reset), then observables
    outcomes = []
    for w in range(4):
        m = dev.measure(w)
        outcomes.append(int(m))
        if m:
            dev.named_op("PauliX", [], [w])
    out["mcm_outcomes"] = np.asarray(outcomes, dtype=np.int64)
    out["mcm_expvals"] = np.asarray([dev.expval(dev.obs_named("PauliZ", w)) for w in (0, 3, 4, n - 1)])
    out["mcm_probs"] = dev.probs([4, 5, n - 1])
    return out


def c5_program(n, depth):
    return workloads.c5_random_depth(n, depth=depth, seed=0)


def run_c5(dev, n, depth=40, state=True):
    apply_ops(dev, c5_program(n, depth))
    out = {"expval": np.float64(dev.expval(dev.obs_named("PauliZ", 0))),
           "expval_zz": np.float64(dev.expval(dev.obs_tensor([dev.obs_named("PauliZ", 0),
                                                              dev.obs_named("PauliX", n - 1)]))),
           "partial_probs": dev.probs(list(range(10)))}
    if state:
        st = dev.state()
        out["state_probe"] = st[probe_indices(n)]
        out["state"] = st
    return out


FIXTURE_KEYS = {
    "c2": ["expval", "grad", "state_probe"],
    "c3": ["probs_probe", "probs_sum", "sample_sha256", "sample_head"],
    "c4": ["expval", "grad", "mcm_outcomes", "mcm_expvals", "mcm_probs"],
    "c5": ["expval", "expval_zz", "partial_probs", "state_probe"],
}


def expval_err(got, ref, norm=1.0):
    """Error of an expectation value relative to max(|ref|, ||O||): <O> of a near-maximally mixed marginal is
    ~1e-4 while the rounding noise of the 2^n-term sum is set by ||O|| (1 for a Pauli word, sum |c_i| for a
    Hamiltonian), so |ref| alone is the wrong yardstick for "1e-12 relative"."""
    return abs(float(got) - float(ref)) / max(abs(float(ref)), float(norm))


def rel_err(a, b):
    """max |a - b| relative to the largest magnitude of the reference b (the tolerance `north_star`
    states is relative: 1e-12)."""
    a, b = np.asarray(a), np.asarray(b)
    scale = float(np.max(np.abs(b))) if b.size else 1.0
    return float(np.max(np.abs(a - b))) / (scale if scale > 0 else 1.0) if b.size else 0.0
