"""The drop-in boundary itself: the plugin .so (C++ vtable of QuantumDevice behind
`B200StateVectorFactory`) driven by the REFERENCE's own runtime C API compiled from its sources
(oracle/_ref/librt_capi.so), i.e. the exact call sequence of a @qjit-compiled program, checked
against the CPU oracle and the reference's seeded golden vectors."""
import numpy as np
import pytest

import rt_driver
from dsl import apply_ops, build_obs, random_circuit, random_op, random_pauli_obs, random_state
from golden import reference_cases as G

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rt_driver.available(), reason="librt_capi / plugin not built")]
TOL = 1e-12


@pytest.fixture
def rt():
    devs = []

    def make(n, **kw):
        d = rt_driver.RuntimeDevice(n, **kw)
        devs.append(d)
        return d
    yield make
    for d in devs:
        d.close()


@pytest.mark.parametrize("n,seed", [(2, 0), (5, 1), (9, 2), (13, 3), (16, 4)])
def test_runtime_random_circuit(rt, oracle_mod, n, seed):
    ops = random_circuit(seed, n, 60)
    init = random_state(seed + 1, n)
    o = oracle_mod.OracleDevice(n)
    d = rt(n)
    assert d.num_qubits() == n
    o.set_state(init, list(range(n)))
    d.set_state(init, list(range(n)))
    apply_ops(o, ops)
    apply_ops(d, ops)
    assert np.max(np.abs(o.state() - d.state())) < TOL
    rng = np.random.default_rng(seed)
    herm = rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4))
    herm = herm + herm.conj().T
    obs_list = [random_pauli_obs(rng, n) for _ in range(4)] + [("N", "Hadamard", 0)]
    obs_list.append(("Ham", list(rng.normal(size=5)), [random_pauli_obs(rng, n) for _ in range(5)]))
    if n >= 2:
        obs_list.append(("Ham", [0.5, 2.0], [("H", herm, [n - 1, 0]), ("N", "PauliZ", 0)]))
    for obs in obs_list:
        ko, kd = build_obs(o, obs), build_obs(d, obs)
        assert abs(o.expval(ko) - d.expval(kd)) < TOL * 10
        assert abs(o.var(ko) - d.var(kd)) < 1e-10
    w = [int(x) for x in rng.permutation(n)[:max(1, n // 2)]]
    assert np.max(np.abs(o.probs(w) - d.probs(w))) < TOL
    assert np.max(np.abs(o.probs() - d.probs())) < TOL


def test_runtime_golden_constants(rt, oracle_mod):
    for case in G.CASES[:12]:
        _, _, n, ops, kind, obs, expected = case
        d = rt_driver.RuntimeDevice(n)
        try:
            apply_ops(d, ops)
            key = build_obs(d, obs)
            got = d.expval(key) if kind == "expval" else d.var(key)
        finally:
            d.close()
        assert np.isclose(got, expected, rtol=1e-5, atol=1e-8), (case[0], got, expected)


def _seeded_program(dev, x, terminal):
    out = []
    for _ in range(4):
        dev.reset(1)  # release + device_init again: the pooled device object is reused
        dev.named_op("RX", [x], [0])
        if dev.measure(0):
            dev.named_op("Hadamard", [], [0])
        out.append(terminal(dev))
    return out


def test_runtime_seeded_mcm_changelog_0_8_0(rt):
    g = G.SEEDED_MCM
    d = rt(1, seed=g["seed"])
    rows = _seeded_program(d, g["x"], lambda dev: dev.probs())
    assert np.allclose(np.array(rows), np.array(g["probs"]))


def test_runtime_seeded_sample_changelog_0_9_0(rt):
    g = G.SEEDED_SAMPLE
    d = rt(1, seed=g["seed"], shots=g["shots"])
    rows = _seeded_program(d, g["x"], lambda dev: dev.sample(g["shots"])[:, 0])
    assert np.array_equal(np.array(rows, dtype=int), np.array(g["samples"]))


@pytest.mark.parametrize("n,seed", [(4, 0), (10, 1), (14, 2)])
def test_runtime_seeded_samples_and_counts(rt, oracle_mod, n, seed):
    ops = random_circuit(seed, n, 30)
    shots = 1000
    o = oracle_mod.OracleDevice(n, seed=99)
    d = rt(n, seed=99, shots=shots)
    apply_ops(o, ops)
    apply_ops(d, ops)
    assert np.array_equal(o.sample(shots), d.sample(shots))
    w = [n - 1, 0]
    assert np.array_equal(o.sample(shots, w), d.sample(shots, w))
    eo, co = o.counts(shots, w)
    ed, cd = d.counts(shots, w)
    assert np.array_equal(eo, ed) and np.array_equal(co, cd)
    assert o.measure(1) == d.measure(1)


def test_runtime_pauli_measure(rt, oracle_mod):
    """__catalyst__qis__PauliMeasure (RuntimeCAPI.cpp:962-993) -> QuantumDevice::PauliMeasure of the plugin,
    incl. the runtime-side `negated` flip."""
    n = 6
    ops = random_circuit(9, n, 30)
    o = oracle_mod.OracleDevice(n, seed=37)
    d = rt(n, seed=37)
    apply_ops(o, ops)
    apply_ops(d, ops)
    for word, wires, neg in [("XZ", [0, 3], False), ("YYX", [5, 1, 2], True), ("Z", [4], False), ("XIZY", [2, 0, 5, 3], True)]:
        assert d.pauli_measure(word, wires, negated=neg) == (o.pauli_measure(word, wires) != neg)
    assert np.max(np.abs(o.state() - d.state())) < 1e-11


DIFF = ["RX", "RY", "RZ", "PhaseShift", "IsingXX", "IsingZZ", "CRX", "CRY", "CRZ", "ControlledPhaseShift",
        "SingleExcitation", "DoubleExcitation", "MultiRZ", "IsingXY", "IsingYY"]


@pytest.mark.parametrize("n,seed", [(4, 0), (9, 1), (13, 2)])
def test_runtime_adjoint_gradient(rt, oracle_mod, n, seed):
    rng = np.random.default_rng(seed)
    ops = [random_op(rng, n, names=[DIFF[rng.integers(len(DIFF))]] if rng.random() < 0.7 else
                     ["Hadamard", "CNOT", "S", "Toffoli", "SWAP"]) for _ in range(25)]
    obs_list = [("N", "PauliZ", 0), ("T", [("N", "PauliX", 1), ("N", "PauliY", 3)]),
                ("Ham", [0.3, -0.8], [("N", "PauliZ", 2), ("N", "Hadamard", n - 1)])]
    o = oracle_mod.OracleDevice(n)
    o.start_tape()
    apply_ops(o, [("Hadamard", [], [w]) for w in range(n)] + ops)
    vo = [o.expval(build_obs(o, ob)) for ob in obs_list]
    o.stop_tape()
    jo = o.gradient()
    d = rt(n)
    d.start_tape()  # toggle_recorder(true) with an active device -> StartTapeRecording
    apply_ops(d, [("Hadamard", [], [w]) for w in range(n)] + ops)
    vd = [d.expval(build_obs(d, ob)) for ob in obs_list]
    d.stop_tape()
    jd = d.gradient(len(obs_list), jo.shape[1])
    assert np.max(np.abs(np.array(vo) - np.array(vd))) < TOL
    assert np.max(np.abs(jo - jd)) < TOL * 10
    if jo.shape[1] >= 2:
        sub = [0, jo.shape[1] - 1]
        assert np.max(np.abs(d.gradient(len(obs_list), jo.shape[1], train=sub) - jo[:, sub])) < TOL * 10
