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
