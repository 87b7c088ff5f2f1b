"""Pin the CPU oracle against every known-answer vector the reference holds for this path
(SURVEY.md 8(c)): analytic constants from the reference's frontend tests and the two seeded
vectors from its release notes."""
import numpy as np
import pytest

from dsl import apply_ops, build_obs
from golden import reference_cases as G


@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_expval_var_constants(oracle_mod, case):
    _, _, n, ops, kind, obs, expected = case
    dev = oracle_mod.OracleDevice(n)
    apply_ops(dev, ops)
    key = build_obs(dev, obs)
    got = dev.expval(key) if kind == "expval" else dev.var(key)
    assert np.isclose(got, expected, rtol=1e-5, atol=1e-8), (got, expected)


@pytest.mark.parametrize("case", G.STATE_CASES, ids=[c[0] for c in G.STATE_CASES])
def test_state_constants(oracle_mod, case):
    _, _, n, ops, expected = case
    dev = oracle_mod.OracleDevice(n)
    apply_ops(dev, ops)
    assert np.allclose(dev.state(), expected)


@pytest.mark.parametrize("case", G.PROBS_CASES, ids=[c[0] for c in G.PROBS_CASES])
def test_probs_constants(oracle_mod, case):
    _, _, n, ops, wires, expected = case
    dev = oracle_mod.OracleDevice(n)
    apply_ops(dev, ops)
    assert np.allclose(dev.probs(wires), expected)


@pytest.mark.parametrize("case", G.SAMPLE_CASES, ids=[c[0] for c in G.SAMPLE_CASES])
def test_sample_deterministic(oracle_mod, case):
    _, _, n, ops, shots, row = case
    dev = oracle_mod.OracleDevice(n)
    apply_ops(dev, ops)
    s = dev.sample(shots)
    assert s.shape == (shots, n)
    assert np.array_equal(s, np.array([row] * shots, dtype=float))


@pytest.mark.parametrize("case", G.COUNTS_CASES, ids=[c[0] for c in G.COUNTS_CASES])
def test_counts_deterministic(oracle_mod, case):
    _, _, n, ops, shots, expected = case
    dev = oracle_mod.OracleDevice(n)
    apply_ops(dev, ops)
    eig, cnt = dev.counts(shots)
    assert np.array_equal(eig, np.arange(1 << n))
    assert np.array_equal(cnt, expected)


def _seeded_program(dev, x, terminal):
    out = []
    for _ in range(4):
        dev.reset(1)
        dev.named_op("RX", [x], [0])
        if dev.measure(0):
            dev.named_op("Hadamard", [], [0])
        out.append(terminal(dev))
    return out


def test_seeded_mcm_changelog_0_8_0(oracle_mod):
    g = G.SEEDED_MCM
    dev = oracle_mod.OracleDevice(1, seed=g["seed"])
    rows = _seeded_program(dev, g["x"], lambda d: d.probs())
    assert np.allclose(np.array(rows), np.array(g["probs"]))


def test_seeded_sample_changelog_0_9_0(oracle_mod):
    g = G.SEEDED_SAMPLE
    dev = oracle_mod.OracleDevice(1, seed=g["seed"])
    rows = _seeded_program(dev, g["x"], lambda d: d.sample(g["shots"])[:, 0])
    assert np.array_equal(np.array(rows, dtype=int), np.array(g["samples"]))


@pytest.mark.parametrize("case", G.DYNAMIC_CASES, ids=[c[0] for c in G.DYNAMIC_CASES])
def test_dynamic_qubit_constants(oracle_mod, case):
    """qml.allocate / deallocate programs of the reference's test_dynamic_qubit_allocation.py."""
    from dsl import DynamicRegister
    _, _, n, prog, wires, want = case
    o = oracle_mod.OracleDevice(n, seed=37)
    r = DynamicRegister(o, n)
    prog(r)
    assert o.n == n and r.ids == list(range(n))
    assert np.allclose(r.probs(wires), want, atol=1e-12)


def test_dynamic_qubit_release_keeps_the_rest(oracle_mod):
    """Releasing a qubit in a basis state leaves the state of the other wires untouched, wherever the
    wire sits, and later wires move down (QubitManager.hpp:124-126)."""
    from dsl import apply_ops, random_circuit
    n = 5
    ops = random_circuit(9, n, 40)
    want = oracle_mod.OracleDevice(n)
    apply_ops(want, ops)
    for at in range(n + 1):                       # the extra wire inserted at every position
        for bit in (0, 1):
            o = oracle_mod.OracleDevice(n + 1)
            if bit:
                o.named_op("PauliX", [], [at])
            shifted = [(op[0], op[1], [w + (w >= at) for w in op[2]]) + tuple(op[3:]) for op in ops]
            shifted = [s if len(s) < 4 else (s[0], s[1], s[2], {**s[3], "ctrl_wires": [w + (w >= at) for w in s[3].get("ctrl_wires", ())]}) for s in shifted]
            apply_ops(o, shifted)
            o.release_qubit(at)
            assert o.n == n
            assert np.max(np.abs(o.state() - want.state())) < 1e-12


@pytest.mark.parametrize("seed", [None, 37, 1234])
@pytest.mark.parametrize("case", G.MCM_CASES, ids=[c[0] for c in G.MCM_CASES])
def test_mcm_identities(oracle_mod, case, seed):
    _, _, n, prog = case
    assert prog(oracle_mod.OracleDevice(n, seed=seed))


@pytest.mark.parametrize("case", G.GRADIENT_CASES, ids=[c[0] for c in G.GRADIENT_CASES])
def test_adjoint_gradient_closed_form(oracle_mod, case):
    _, _, n, ops, obs, want = case
    o = oracle_mod.OracleDevice(n)
    o.start_tape()
    apply_ops(o, ops)
    o.expval(build_obs(o, obs))
    o.stop_tape()
    assert np.allclose(o.gradient(), [want], rtol=0, atol=1e-12)
