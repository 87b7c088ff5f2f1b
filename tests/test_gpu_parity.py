"""Parity of the CUDA device (through the C ABI of include/b200sv.h) against the CPU oracle and
against the reference's own known-answer vectors.  Tolerance: 1e-12 on amplitudes, expectation
values and gradients (BASELINE.json north_star); bit-exact for seeded mid-circuit measurement
outcomes and sample bitstrings."""
import numpy as np
import pytest

from dsl import (apply_ops, assert_close, build_obs, random_circuit, random_op, random_pauli_obs, random_state)
from golden import reference_cases as G

pytestmark = pytest.mark.gpu
TOL = 1e-12


@pytest.fixture(scope="module")
def gpu():
    from catalyst_b200 import capi
    return capi


# ---- reference known-answer vectors ------------------------------------------------------------
@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_expval_var_constants(gpu, case):
    _, _, n, ops, kind, obs, expected = case
    dev = gpu.B200Device(n)
    apply_ops(dev, ops)
    key = build_obs(dev, obs)
    got = dev.expval(key) if kind == "expval" else dev.var(key)
    assert np.isclose(got, expected, rtol=1e-5, atol=1e-8), (got, expected)


@pytest.mark.parametrize("case", G.STATE_CASES, ids=[c[0] for c in G.STATE_CASES])
def test_state_constants(gpu, case):
    _, _, n, ops, expected = case
    dev = gpu.B200Device(n)
    apply_ops(dev, ops)
    assert np.allclose(dev.state(), expected)


@pytest.mark.parametrize("case", G.PROBS_CASES, ids=[c[0] for c in G.PROBS_CASES])
def test_probs_constants(gpu, case):
    _, _, n, ops, wires, expected = case
    dev = gpu.B200Device(n)
    apply_ops(dev, ops)
    assert np.allclose(dev.probs(wires), expected)


@pytest.mark.parametrize("case", G.SAMPLE_CASES, ids=[c[0] for c in G.SAMPLE_CASES])
def test_sample_deterministic(gpu, case):
    _, _, n, ops, shots, row = case
    dev = gpu.B200Device(n)
    apply_ops(dev, ops)
    s = dev.sample(shots)
    assert np.array_equal(s, np.array([row] * shots, dtype=float))


@pytest.mark.parametrize("case", G.COUNTS_CASES, ids=[c[0] for c in G.COUNTS_CASES])
def test_counts_deterministic(gpu, case):
    _, _, n, ops, shots, expected = case
    dev = gpu.B200Device(n)
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


def test_seeded_mcm_changelog_0_8_0(gpu):
    g = G.SEEDED_MCM
    dev = gpu.B200Device(1, seed=g["seed"])
    rows = _seeded_program(dev, g["x"], lambda d: d.probs())
    assert np.allclose(np.array(rows), np.array(g["probs"]))


def test_seeded_sample_changelog_0_9_0(gpu):
    g = G.SEEDED_SAMPLE
    dev = gpu.B200Device(1, seed=g["seed"])
    rows = _seeded_program(dev, g["x"], lambda d: d.sample(g["shots"])[:, 0])
    assert np.array_equal(np.array(rows, dtype=int), np.array(g["samples"]))


# ---- random circuits vs the oracle --------------------------------------------------------------
CONFIGS = [dict(), dict(fuse=False), dict(tile_bits=4, min_row_bits=2), dict(tile_bits=6, min_row_bits=0),
           dict(tile_bits=12, min_row_bits=7),
           # rows shorter than the tile's scattered-bit limit allows are clamped, not an error at the first pass
           dict(tile_bits=11, min_row_bits=0, jit="always")]


@pytest.mark.parametrize("cfg", range(len(CONFIGS)))
@pytest.mark.parametrize("n,seed", [(1, 0), (2, 1), (3, 2), (5, 3), (7, 4), (9, 5), (11, 6), (12, 7),
                                    (13, 8), (14, 9), (16, 10)])
def test_random_circuit_state(gpu, oracle_mod, n, seed, cfg):
    ops = random_circuit(seed, n, 60)
    init = random_state(seed + 500, n)
    o = oracle_mod.OracleDevice(n)
    d = gpu.B200Device(n, **CONFIGS[cfg])
    o.set_state(init, list(range(n)))
    d.set_state(init, list(range(n)))
    apply_ops(o, ops)
    apply_ops(d, ops)
    a, b = o.state(), d.state()
    assert np.max(np.abs(a - b)) < TOL
    assert abs(np.linalg.norm(b) - 1) < 1e-12


@pytest.mark.parametrize("n", [18, 20])
def test_random_circuit_state_large(gpu, oracle_mod, n):
    ops = random_circuit(n, n, 80)
    o = oracle_mod.OracleDevice(n)
    d = gpu.B200Device(n)
    apply_ops(o, [("Hadamard", [], [w]) for w in range(n)] + ops)
    apply_ops(d, [("Hadamard", [], [w]) for w in range(n)] + ops)
    assert np.max(np.abs(o.state() - d.state())) < TOL


@pytest.mark.parametrize("k", [2, 3, 4, 5, 6, 8])
def test_matrix_operation_k_wires(gpu, oracle_mod, k):
    from dsl import random_unitary
    n = 12
    rng = np.random.default_rng(k)
    init = random_state(k, n)
    o, d = oracle_mod.OracleDevice(n), gpu.B200Device(n)
    o.set_state(init, list(range(n)))
    d.set_state(init, list(range(n)))
    for trial in range(3):
        perm = [int(w) for w in rng.permutation(n)]
        U = random_unitary(rng, k)
        kw = dict(inverse=bool(trial & 1))
        if trial == 2:
            kw.update(ctrl_wires=perm[k:k + 2], ctrl_values=[True, False])
        o.matrix_op(U, perm[:k], **kw)
        d.matrix_op(U, perm[:k], **kw)
    assert np.max(np.abs(o.state() - d.state())) < TOL


def test_state_prep_and_wire_order(gpu, oracle_mod):
    d = gpu.B200Device(3)
    d.named_op("PauliX", [], [0])
    assert np.argmax(np.abs(d.state())) == 4
    assert np.allclose(d.probs([2, 0]), [0, 1, 0, 0])
    sub = random_state(3, 2)
    o = oracle_mod.OracleDevice(5)
    d = gpu.B200Device(5)
    o.set_state(sub, [3, 1])
    d.set_state(sub, [3, 1])
    assert np.allclose(o.state(), d.state(), atol=0)
    o.set_basis_state([1, 0, 1], [4, 2, 0])
    d.set_basis_state([1, 0, 1], [4, 2, 0])
    assert np.array_equal(o.state(), d.state())


# ---- measurements ------------------------------------------------------------------------------
def _prepared(gpu, oracle_mod, n, seed, depth=40, **kw):
    ops = random_circuit(seed, n, depth)
    o = oracle_mod.OracleDevice(n, seed=kw.pop("rng_seed", None))
    d = gpu.B200Device(n, **kw)
    apply_ops(o, ops)
    apply_ops(d, ops)
    return o, d


@pytest.mark.parametrize("n,seed", [(3, 0), (6, 1), (10, 2), (13, 3), (17, 4)])
def test_expval_var_random_observables(gpu, oracle_mod, n, seed):
    o, d = _prepared(gpu, oracle_mod, n, seed)
    rng = np.random.default_rng(seed)
    herm = rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4))
    herm = herm + herm.conj().T
    obs_list = [("N", "PauliZ", 0), ("N", "PauliX", n - 1), ("N", "PauliY", n // 2),
                ("N", "Hadamard", 0), ("N", "Identity", 0)]
    obs_list += [random_pauli_obs(rng, n) for _ in range(6)]
    ham_terms = [random_pauli_obs(rng, n) for _ in range(12)]
    obs_list.append(("Ham", list(rng.normal(size=12)), ham_terms))
    if n >= 3:
        obs_list.append(("H", herm, [n - 1, 0]))
        obs_list.append(("T", [("H", herm, [0, 2]), ("N", "PauliX", 1)]))
        obs_list.append(("Ham", [0.3, -1.2, 0.5], [("H", herm, [1, 0]), ("N", "Hadamard", 2),
                                                   ("T", [("N", "PauliZ", 0), ("N", "Hadamard", 1)])]))
        # all-Z Hamiltonian (one fused read pass)
        obs_list.append(("Ham", list(rng.normal(size=n - 1)),
                         [("T", [("N", "PauliZ", i), ("N", "PauliZ", i + 1)]) for i in range(n - 1)]))
    for obs in obs_list:
        ko, kd = build_obs(o, obs), build_obs(d, obs)
        eo, ed = o.expval(ko), d.expval(kd)
        assert abs(eo - ed) < TOL * max(1.0, abs(eo)), (obs, eo, ed)
        vo, vd = o.var(ko), d.var(kd)
        assert_close(vd, vo, "Var (random observables)")


@pytest.mark.parametrize("n,seed", [(1, 0), (4, 1), (9, 2), (12, 3), (15, 4), (19, 5)])
def test_probs_partial(gpu, oracle_mod, n, seed):
    o, d = _prepared(gpu, oracle_mod, n, seed)
    rng = np.random.default_rng(seed)
    assert np.max(np.abs(o.probs() - d.probs())) < TOL
    for _ in range(6):
        k = int(rng.integers(1, n + 1))
        wires = [int(w) for w in rng.permutation(n)[:k]]
        po, pd = o.probs(wires), d.probs(wires)
        assert np.max(np.abs(po - pd)) < TOL, wires
    assert abs(d.probs().sum() - 1) < 1e-12


@pytest.mark.parametrize("n,seed", [(2, 0), (5, 1), (8, 2), (11, 3), (14, 4)])
def test_seeded_samples_bit_exact(gpu, oracle_mod, n, seed):
    ops = random_circuit(seed, n, 30)
    o = oracle_mod.OracleDevice(n, seed=1234 + seed)
    d = gpu.B200Device(n, seed=1234 + seed)
    apply_ops(o, ops)
    apply_ops(d, ops)
    assert np.array_equal(o.sample(2000), d.sample(2000))
    wires = [n - 1, 0] if n > 1 else [0]
    assert np.array_equal(o.sample(500, wires), d.sample(500, wires))
    eo, co = o.counts(1000, wires)
    ed, cd = d.counts(1000, wires)
    assert np.array_equal(eo, ed) and np.array_equal(co, cd)
    assert o.gen_peek() == d.gen_peek()  # same position in the runtime generator's stream


def test_samples_sparse_distributions(gpu, oracle_mod):
    # distributions with many exact zeros / ties stress the prefix-sum restatement of the alias table
    for n, prep in [(6, [("Hadamard", [], [0])]),
                    (10, [("Hadamard", [], [0])] + [("CNOT", [], [0, w]) for w in range(1, 10)]),
                    (12, [("Hadamard", [], [w]) for w in (1, 5, 7)] + [("RY", [0.3], [2])]),
                    (9, [("RY", [2.0], [w]) for w in range(9)])]:
        o = oracle_mod.OracleDevice(n, seed=7)
        d = gpu.B200Device(n, seed=7)
        apply_ops(o, prep)
        apply_ops(d, prep)
        assert np.array_equal(o.sample(3000), d.sample(3000)), (n, prep)


@pytest.mark.parametrize("n,seed", [(3, 0), (8, 1), (12, 2), (16, 3)])
def test_seeded_mid_circuit_measurements(gpu, oracle_mod, n, seed):
    rng = np.random.default_rng(seed)
    o = oracle_mod.OracleDevice(n, seed=37)
    d = gpu.B200Device(n, seed=37)
    outcomes_o, outcomes_d = [], []
    for layer in range(6):
        ops = [random_op(rng, n) for _ in range(10)]
        apply_ops(o, ops)
        apply_ops(d, ops)
        w = int(rng.integers(n))
        outcomes_o.append(o.measure(w))
        outcomes_d.append(d.measure(w))
    assert outcomes_o == outcomes_d
    assert np.max(np.abs(o.state() - d.state())) < TOL
    # postselection
    w = 0
    p = o.probs([w])
    ps = int(p[1] > p[0])
    assert o.measure(w, ps) == d.measure(w, ps) == bool(ps)
    assert_close(d.state(), o.state(), "state after measurement / relabelling")
    with pytest.raises(RuntimeError):
        d.measure(w, 1 - ps)  # probability of the other branch is now exactly 0


# ---- PauliMeasure (QuantumDevice.hpp:531): outcomes bit-exact under the seeded stream, states 1e-12 ----
def _random_word(rng, n):
    k = int(rng.integers(1, min(n, 5) + 1))
    wires = [int(w) for w in rng.permutation(n)[:k]]
    word = "".join(rng.choice(list("IXYZ"), size=k, p=[0.1, 0.3, 0.3, 0.3]))
    return word, wires


@pytest.mark.parametrize("cfg", [dict(), dict(jit="always"), dict(tile_bits=5, min_row_bits=2)])
@pytest.mark.parametrize("n,seed", [(1, 0), (4, 1), (9, 2), (13, 3), (16, 4)])
def test_seeded_pauli_measurements(gpu, oracle_mod, n, seed, cfg):
    rng = np.random.default_rng(seed)
    o = oracle_mod.OracleDevice(n, seed=37)
    d = gpu.B200Device(n, seed=37, **cfg)
    out_o, out_d = [], []
    for layer in range(6):
        ops = [random_op(rng, n) for _ in range(8)]
        apply_ops(o, ops)
        apply_ops(d, ops)
        word, wires = _random_word(rng, n)
        out_o.append(o.pauli_measure(word, wires))
        out_d.append(d.pauli_measure(word, wires))
        if layer % 2:   # interleave with ordinary measurements on the same stream
            w = int(rng.integers(n))
            out_o.append(o.measure(w))
            out_d.append(d.measure(w))
    assert out_o == out_d
    assert_close(d.state(), o.state(), "state after measurement / relabelling")
    assert abs(np.linalg.norm(d.state()) - 1) < 1e-12


def test_pauli_measure_identities_and_errors(gpu):
    d = gpu.B200Device(3, seed=5)
    apply_ops(d, [("Hadamard", [], [0]), ("CNOT", [], [0, 1]), ("CNOT", [], [1, 2])])   # GHZ
    ref = d.state()
    assert [d.pauli_measure("XXX", [0, 1, 2]), d.pauli_measure("ZZ", [0, 1]), d.pauli_measure("ZIZ", [2, 1, 0]),
            d.pauli_measure("II", [0, 2]), d.pauli_measure("YYX", [0, 1, 2])] == [False, False, False, False, True]
    assert np.max(np.abs(d.state() - ref)) < TOL
    r = d.pauli_measure("Z", [1])                      # collapses the GHZ state
    want = np.zeros(8, dtype=complex); want[7 if r else 0] = 1
    assert np.max(np.abs(d.state() - want)) < TOL
    assert d.pauli_measure("ZZ", [0, 2], postselect=0) is False
    with pytest.raises(RuntimeError):   # |00>: the probability of eigenvalue -1 is exactly 0
        gpu.B200Device(2).pauli_measure("ZZ", [0, 1], postselect=1)
    with pytest.raises(RuntimeError):
        d.pauli_measure("XX", [0])
    with pytest.raises(RuntimeError):
        d.pauli_measure("A", [0])


# ---- finite shots: sampled Expval / Var / Probs.  The device rotates in place and
# back, the oracle rotates a copy; both draw the same sample indices from the same stream, so the
# estimates agree to rounding of a mean of +-1 (the states to 1e-12).
@pytest.mark.parametrize("cfg", [dict(), dict(jit="always")])
@pytest.mark.parametrize("n,seed", [(2, 0), (6, 1), (11, 2), (15, 3)])
def test_shots_expval_var_probs(gpu, oracle_mod, n, seed, cfg):
    rng = np.random.default_rng(seed)
    ops = random_circuit(seed, n, 40)
    shots = 5000
    o = oracle_mod.OracleDevice(n, seed=37)
    d = gpu.B200Device(n, seed=37, **cfg)
    apply_ops(o, ops); apply_ops(d, ops)
    o.set_shots(shots); d.set_shots(shots)
    names = ["PauliX", "PauliY", "PauliZ", "Hadamard", "Identity"]
    exact_matches, total = 0, 0
    for trial in range(4):
        terms, coeffs = [], []
        for _ in range(int(rng.integers(1, 4))):
            k = int(rng.integers(1, min(n, 3) + 1))
            ws = [int(w) for w in rng.permutation(n)[:k]]
            terms.append(("T", [("N", names[int(rng.integers(5))], w) for w in ws]))
            coeffs.append(float(rng.normal()))
        ham = ("Ham", coeffs, terms)
        eo, ed = o.expval(build_obs(o, ham)), d.expval(build_obs(d, ham))
        vo, vd = o.var(build_obs(o, terms[0])), d.var(build_obs(d, terms[0]))
        # Same stream, same sampler => same estimate, EXCEPT where a marginal is uniform in exact
        # arithmetic (N * p[i] == 1 up to rounding): the alias table's LIFO pairing then depends on the
        # last bit of p and the two runs draw different, equally valid samples.  Those cases must
        # still agree with the analytic value.
        o.set_shots(0)
        exact_e, exact_v = o.expval(build_obs(o, ham)), o.var(build_obs(o, terms[0]))
        o.set_shots(shots)
        sigma = 5 * sum(abs(c) for c in coeffs) / np.sqrt(shots)
        total += 2
        exact_matches += int(abs(eo - ed) < 1e-12) + int(abs(vo - vd) < 1e-12)
        assert abs(eo - ed) < 1e-12 or (abs(eo - exact_e) < sigma and abs(ed - exact_e) < sigma), (eo, ed, exact_e)
        assert abs(vo - vd) < 1e-12 or (abs(vo - exact_v) < 0.1 and abs(vd - exact_v) < 0.1), (vo, vd, exact_v)
        w = [int(x) for x in rng.permutation(n)[:min(n, 3)]]
        po, pd = o.probs(w), d.probs(w)
        assert np.array_equal(po, pd) or np.max(np.abs(po - pd)) < 0.05
    assert exact_matches >= total // 2, (exact_matches, total)
    # sampled expectation values leave the state where it was
    assert np.max(np.abs(o.state() - d.state())) < TOL
    d.set_shots(0); o.set_shots(0)
    ham = ("Ham", [0.5, -1.5], [("N", "PauliX", 0), ("T", [("N", "PauliZ", 0), ("N", "PauliY", n - 1)])] if n > 1 else
           [("N", "PauliX", 0), ("N", "PauliY", 0)])
    exact = d.expval(build_obs(d, ham))
    d.set_shots(200000)
    assert abs(d.expval(build_obs(d, ham)) - exact) < 5 * 2.0 / np.sqrt(200000)
    # Hermitian observables are sampled in their eigenbasis (host-side diagonalisation): exercised below
    d.set_shots(0)


@pytest.mark.parametrize("n,seed", [(3, 0), (7, 1), (12, 2)])
def test_shots_hermitian_observables(gpu, oracle_mod, n, seed):
    """Finite-shot Expval / Var of Hermitian observables (alone and inside tensor products / Hamiltonians): the
    matrix is diagonalised on the host (eigenvalues ascending), the state rotated by V^dagger, the wires sampled
    and the eigenvalues averaged.  Same stream on device and oracle => the same estimate (or, where a marginal is
    exactly uniform, two estimates within 5 sigma of the analytic value); the state is left where it was."""
    rng = np.random.default_rng(seed)
    ops = random_circuit(seed + 40, n, 40)
    shots = 20000
    o, d = oracle_mod.OracleDevice(n, seed=37), gpu.B200Device(n, seed=37)
    apply_ops(o, ops); apply_ops(d, ops)

    def herm(k):
        m = rng.normal(size=(1 << k, 1 << k)) + 1j * rng.normal(size=(1 << k, 1 << k))
        return m + m.conj().T
    perm = [int(w) for w in rng.permutation(n)]
    cases = [("H", herm(1), [perm[0]]), ("H", herm(2), perm[:2]),
             ("T", [("H", herm(2), [perm[1], perm[0]]), ("N", "PauliX", perm[2])]),
             ("Ham", [0.7, -0.3], [("H", herm(1), [perm[2]]), ("T", [("N", "PauliZ", perm[0]), ("H", herm(1), [perm[1]])])])]
    same = 0
    for obs in cases:
        exact_e = o.expval(build_obs(o, obs))
        spread = 4.0 * (1 << 2)                                   # loose bound on the eigenvalue range
        o.set_shots(shots); d.set_shots(shots)
        eo, ed = o.expval(build_obs(o, obs)), d.expval(build_obs(d, obs))
        if obs[0] != "Ham":
            vo, vd = o.var(build_obs(o, obs)), d.var(build_obs(d, obs))
            assert abs(vo - vd) < 1e-9 or abs(vo - vd) < 0.2 * max(1.0, abs(vo))
        o.set_shots(0); d.set_shots(0)
        same += int(abs(eo - ed) < 1e-10)
        assert abs(eo - ed) < 1e-10 or (abs(eo - exact_e) < 5 * spread / np.sqrt(shots) and
                                          abs(ed - exact_e) < 5 * spread / np.sqrt(shots)), (eo, ed, exact_e)
    assert same >= 2
    assert_close(d.state(), o.state(), "state after measurement / relabelling")


# ---- adjoint gradient ----------------------------------------------------------------------------
DIFF_GATES = ["RX", "RY", "RZ", "PhaseShift", "IsingXX", "IsingYY", "IsingXY", "IsingZZ",
              "ControlledPhaseShift", "CRX", "CRY", "CRZ", "SingleExcitation", "DoubleExcitation",
              "MultiRZ", "PauliRot", "GlobalPhase", "PSWAP"]


@pytest.mark.parametrize("n,seed", [(4, 0), (5, 1), (8, 2), (12, 3), (14, 4)])
def test_adjoint_gradient(gpu, oracle_mod, n, seed):
    rng = np.random.default_rng(seed)
    ops = []
    for _ in range(30):
        if rng.random() < 0.7:
            ops.append(random_op(rng, n, names=[DIFF_GATES[rng.integers(len(DIFF_GATES))]]))
        else:
            ops.append(random_op(rng, n, names=["Hadamard", "CNOT", "S", "QubitUnitary", "Toffoli", "SWAP"]))
    obs_list = [("N", "PauliZ", 0), ("T", [("N", "PauliX", 1), ("N", "PauliY", 3)]),
                ("Ham", [0.3, -0.8], [("N", "PauliZ", 2), ("H", [[1, 2], [2, -1]], [n - 1])])]
    devs = [oracle_mod.OracleDevice(n), gpu.B200Device(n)]
    jacs = []
    for dev in devs:
        dev.set_state(random_state(99, n), list(range(n)))
        dev.start_tape()
        apply_ops(dev, ops)
        vals = [dev.expval(build_obs(dev, ob)) for ob in obs_list]
        dev.stop_tape()
        jacs.append((np.array(vals), dev.gradient()))
    assert np.max(np.abs(jacs[0][0] - jacs[1][0])) < TOL
    assert jacs[0][1].shape == jacs[1][1].shape
    assert np.max(np.abs(jacs[0][1] - jacs[1][1])) < TOL * 10
    P = jacs[0][1].shape[1]
    if P >= 3:
        sub = [0, P // 2, P - 1]
        assert np.max(np.abs(devs[1].gradient(sub) - jacs[0][1][:, sub])) < TOL * 10


def test_gradient_rejects_var_and_multiparam(gpu):
    d = gpu.B200Device(2)
    d.start_tape()
    d.named_op("RX", [0.3], [0])
    d.var(d.obs_named("PauliZ", 0))
    d.stop_tape()
    with pytest.raises(RuntimeError):
        d.gradient()
    d = gpu.B200Device(2)
    d.start_tape()
    d.named_op("Rot", [0.3, 0.1, 0.2], [0])
    d.expval(d.obs_named("PauliZ", 0))
    d.stop_tape()
    with pytest.raises(RuntimeError):
        d.gradient()


# ---- error behaviour + lifecycle -------------------------------------------------------------------
def test_errors_and_reuse(gpu):
    d = gpu.B200Device(3)
    with pytest.raises(RuntimeError):
        d.named_op("NotAGate", [], [0])
    with pytest.raises(RuntimeError):
        d.named_op("RX", [], [0])
    with pytest.raises(RuntimeError):
        d.named_op("CNOT", [], [0, 0])
    with pytest.raises(RuntimeError):
        d.named_op("PauliX", [], [3])
    with pytest.raises(RuntimeError):
        d.expval(17)
    d.named_op("PauliX", [], [1])
    d.release_qubits()
    d.allocate_qubits(2)  # pooled-device reuse (ExecutionContext.hpp:355-361): clean |00>
    assert np.array_equal(d.state(), [1, 0, 0, 0])
    d.allocate_qubit()
    d.named_op("PauliX", [], [2])
    assert np.argmax(np.abs(d.state())) == 1 and d.num_qubits() == 3


# ---- size-independent properties at a size the oracle does not reach in seconds --------------------
def test_large_roundtrip_properties(gpu):
    n = 26
    d = gpu.B200Device(n)
    ops = [("Hadamard", [], [w]) for w in range(n)] + random_circuit(5, n, 120)
    apply_ops(d, ops)
    inv = []
    for op in reversed(ops):
        kw = dict(op[3]) if len(op) > 3 else {}
        kw["inverse"] = not kw.get("inverse", False)
        inv.append((op[0], op[1], op[2], kw))
    p = d.probs([0, 1, 2])
    assert abs(p.sum() - 1) < 1e-12
    apply_ops(d, inv)
    # back to |0...0>: probability of the all-zero string is 1
    assert abs(d.probs([n - 1])[0] - 1) < 1e-10
    assert abs(d.expval(d.obs_named("PauliZ", 0)) - 1) < 1e-10


# ---- SWAP gates as wire relabelling ---------------------------------------------------------------
@pytest.mark.parametrize("cfg", [dict(), dict(jit="always"), dict(virtual_swap=False)])
@pytest.mark.parametrize("n,seed", [(2, 0), (5, 1), (9, 2), (12, 3), (15, 4)])
def test_swap_relabelling(gpu, oracle_mod, n, seed, cfg):
    """Uncontrolled SWAPs move no data (the wires trade index bits); every way of looking at the state
    must still see the oracle's wire order: State, Probs (all wires: permuted on the way out), partial
    probs, expval, seeded samples on all wires (bit-exact: the alias table is built over the canonical
    bin order), Measure, and gates applied after the swaps."""
    rng = np.random.default_rng(seed)
    ops = []
    for layer in range(6):
        ops += [random_op(rng, n) for _ in range(6)]
        for _ in range(3):
            a, b = (int(x) for x in rng.permutation(n)[:2])
            ops.append(("SWAP", [], [a, b]))
    o = oracle_mod.OracleDevice(n, seed=37)
    d = gpu.B200Device(n, seed=37, **cfg)
    init = random_state(seed + 77, n)
    o.set_state(init, list(range(n))); d.set_state(init, list(range(n)))
    apply_ops(o, ops); apply_ops(d, ops)
    s = d.stats()
    assert (s.relabelled_swaps >= 18) if cfg.get("virtual_swap", True) else (s.relabelled_swaps == 0)
    assert np.max(np.abs(o.probs() - d.probs())) < TOL                       # all wires, permuted output
    w = [int(x) for x in rng.permutation(n)[:min(n, 3)]]
    assert np.max(np.abs(o.probs(w) - d.probs(w))) < TOL
    obs = random_pauli_obs(rng, n)
    assert abs(o.expval(build_obs(o, obs)) - d.expval(build_obs(d, obs))) < TOL
    assert np.array_equal(o.sample(2000), d.sample(2000))
    assert o.measure(n - 1) == d.measure(n - 1)
    more = [random_op(rng, n) for _ in range(10)]
    apply_ops(o, more); apply_ops(d, more)
    assert_close(d.state(), o.state(), "state after measurement / relabelling")   # State restores the order
    # after State the relabelling is gone and everything still agrees
    assert_close(d.probs(), o.probs(), "probs after relabelling")


def test_swap_relabelling_with_tape_and_dynamic_qubits(gpu, oracle_mod):
    n = 6
    ops = [("RX", [0.3], [0]), ("SWAP", [], [0, 4]), ("RY", [0.7], [4]), ("CNOT", [], [4, 2]), ("SWAP", [], [2, 5]),
           ("RZ", [1.1], [5]), ("IsingXX", [0.4], [1, 5])]
    o, d = oracle_mod.OracleDevice(n), gpu.B200Device(n)
    pre = [("Hadamard", [], [1]), ("SWAP", [], [1, 3]), ("CNOT", [], [3, 0])]      # relabelled before the tape starts
    apply_ops(o, pre); apply_ops(d, pre)
    o.start_tape(); d.start_tape()
    apply_ops(o, ops); apply_ops(d, ops)
    obs = ("Ham", [0.5, -1.0], [("N", "PauliZ", 5), ("T", [("N", "PauliX", 0), ("N", "PauliY", 3)])])
    eo, ed = o.expval(build_obs(o, obs)), d.expval(build_obs(d, obs))
    o.stop_tape(); d.stop_tape()
    assert abs(eo - ed) < TOL
    assert_close(d.gradient(), o.gradient(), "adjoint gradient (closed-form cases)")
    # dynamic allocation on a relabelled register
    d2 = gpu.B200Device(3)
    apply_ops(d2, [("Hadamard", [], [0]), ("SWAP", [], [0, 2]), ("CNOT", [], [2, 1])])
    d2.allocate_qubit()
    apply_ops(d2, [("CNOT", [], [1, 3])])
    o2 = oracle_mod.OracleDevice(4)
    apply_ops(o2, [("Hadamard", [], [0]), ("SWAP", [], [0, 2]), ("CNOT", [], [2, 1]), ("CNOT", [], [1, 3])])
    assert np.max(np.abs(o2.state() - d2.state())) < TOL


def test_many_large_matrix_operations_do_not_wrap_the_arena(gpu, oracle_mod):
    """Five back-to-back 10-wire QubitUnitary ops (16 MiB each) exceed the 64 MiB device matrix arena: the pass
    must be closed before the arena wraps, not overwrite the first matrix."""
    from dsl import random_unitary
    n = 11
    rng = np.random.default_rng(5)
    mats = [random_unitary(rng, 10) for _ in range(5)]
    o, d = oracle_mod.OracleDevice(n), gpu.B200Device(n)
    for dev in (o, d):
        dev.named_op("Hadamard", [], [0])
        for i, m in enumerate(mats):
            wires = [int(w) for w in np.roll(np.arange(n), i)[:10]]
            dev.matrix_op(m, wires)
    assert np.max(np.abs(o.state() - d.state())) < 1e-12


def test_stats_json_and_fp64_peak(gpu):
    d = gpu.B200Device(12, jit="always")
    for w in range(12):
        d.named_op("RX", [0.1 * w], [w])
    d.synchronize()
    js = d.stats_json()
    assert js["num_qubits"] == 12 and js["num_gates"] == 12 and js["tile_passes"] >= 1
    assert js["specialised_passes"] == js["tile_passes"] and js["fp64_instructions"] > 0
    rate = d.fp64_peak()
    assert 1e12 < rate < 1e14, rate       # B200: 148 SMs x 64 FP64 lanes x ~1.9 GHz = 1.8e13 DFMA/s


@pytest.mark.parametrize("kernel", ["Local", "NonZeroRandom"])
@pytest.mark.parametrize("n,seed", [(4, 0), (9, 1), (14, 2)])
def test_mcmc_sampling(gpu, oracle_mod, n, seed, kernel):
    """Device kwargs mcmc / num_burnin / kernel_name (frontend/catalyst/from_plxpr/from_plxpr.py:194): Sample and
    Counts draw from a Metropolis chain.  PARITY UNPINNED against lightning.qubit (its kernels seed themselves from
    std::random_device); pinned here: device == CPU restatement bit for bit under a seed, wires relabelled by SWAPs
    included, and the chain's frequencies converge to the analytic probabilities."""
    ops = random_circuit(seed + 70, n, 40) + [("SWAP", [], [0, n - 1])]
    shots = 4000
    o = oracle_mod.OracleDevice(n, seed=37)
    d = gpu.B200Device(n, seed=37, mcmc=True, num_burnin=200, kernel_name=kernel)
    o.set_mcmc(True, 200, kernel)
    apply_ops(o, ops); apply_ops(d, ops)
    assert np.array_equal(o.sample(shots), d.sample(shots))
    w = [n - 1, 0, n // 2]
    assert np.array_equal(o.sample(shots, w), d.sample(shots, w))
    eo, co = o.counts(shots, w[:2])
    ed, cd = d.counts(shots, w[:2])
    assert np.array_equal(co, cd) and np.array_equal(eo, ed)
    assert_close(d.state(), o.state(), "state after measurement / relabelling")
    if n == 4:
        big = d.sample(200000)
        idx = (big.astype(np.int64) * (1 << np.arange(n - 1, -1, -1))).sum(axis=1)
        freq = np.bincount(idx, minlength=1 << n) / len(idx)
        d2 = gpu.B200Device(n, seed=1)
        apply_ops(d2, ops)
        assert np.max(np.abs(freq - d2.probs())) < 0.02


def test_mcmc_kwargs_are_validated(gpu):
    with pytest.raises(RuntimeError, match="kernel_name"):
        gpu.B200Device(2, mcmc=True, kernel_name="Gibbs")


# ---- tile-staged sums of Pauli words (b200k_pauli_words_tile), called directly through the C ABI -----------------
@pytest.mark.parametrize("n,c,high", [(13, 4, [5, 6, 8, 9, 10, 11, 12]), (9, 3, [3, 4, 5, 6, 7, 8]), (12, 11, []),
                                      (4, 2, [2, 3])])
def test_pauli_words_tile_entry_point(gpu, n, c, high):
    """out = sum_w coef_w P_w in (written, and accumulated on a second call) and <bra| sum_w coef_w P_w |in> against
    numpy, with real, imaginary and general complex coefficients, Z factors inside and outside the tile."""
    import ctypes as C
    import torch
    L = gpu.lib()

    class Hdr(C.Structure):
        _fields_ = [("n_rounds", C.c_int32), ("n_ops", C.c_int32), ("t", C.c_int32), ("c", C.c_int32),
                    ("nhigh", C.c_int32), ("pad", C.c_int32), ("high_pos", C.c_uint8 * 16), ("rank_base", C.c_uint64),
                    ("ntiles", C.c_uint64)]

    class Word(C.Structure):
        _fields_ = [("xin", C.c_uint32), ("zin", C.c_uint32), ("zo", C.c_uint64), ("cr", C.c_double), ("ci", C.c_double),
                    ("ny", C.c_int32), ("pad", C.c_int32)]

    L.b200k_pauli_words_tile.restype = C.c_int
    L.b200k_pauli_words_tile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(Hdr), C.POINTER(Word),
                                         C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.b200k_reduce_scratch_doubles.restype = C.c_int64
    t = c + len(high)
    assert t == min(11, n)
    loc = {p: p for p in range(c)}
    loc.update({p: c + i for i, p in enumerate(high)})
    rng = np.random.default_rng(100 + n)
    da, db = gpu.B200Device(n, seed=1, virtual_swap=False), gpu.B200Device(n, seed=2, virtual_swap=False)
    apply_ops(da, random_circuit(21, n, 30))
    apply_ops(db, random_circuit(22, n, 30))
    va, vb = da.state(), db.state()  # wire 0 = most significant index bit: index i <-> physical amplitude i
    pa, pb = da.device_pointer(), db.device_pointer()
    H = Hdr(t=t, c=c, nhigh=len(high), rank_base=0, ntiles=1 << (n - t))
    for i, p in enumerate(high):
        H.high_pos[i] = p
    nw = 40
    words = (Word * nw)()
    idx = np.arange(1 << n, dtype=np.uint64)
    want = np.zeros(1 << n, dtype=np.complex128)
    inside = sorted(loc)
    for w in range(nw):
        xbits = [int(b) for b in rng.choice(inside, size=int(rng.integers(0, min(4, t) + 1)), replace=False)]
        zbits = [int(b) for b in rng.choice(n, size=int(rng.integers(0, n + 1)), replace=False)]
        x = sum(1 << b for b in xbits)
        z = sum(1 << b for b in zbits)
        ny = bin(x & z).count("1")
        kind = w % 3
        coef = complex(rng.normal(), 0.0) if kind == 0 else complex(0.0, rng.normal()) if kind == 1 else \
            complex(rng.normal(), rng.normal())
        words[w] = Word(xin=sum(1 << loc[b] for b in xbits), zin=sum(1 << loc[b] for b in zbits if b in loc),
                        zo=sum(1 << b for b in zbits if b not in loc), cr=coef.real, ci=coef.imag, ny=ny)
        j = idx ^ np.uint64(x)
        par = np.zeros(1 << n, dtype=np.int64)
        for b in zbits:
            par ^= ((j >> np.uint64(b)) & np.uint64(1)).astype(np.int64)
        want += coef * (1j ** ny) * (1 - 2 * par) * va[j.astype(np.int64)]
    torch.cuda.set_device(0)
    out = torch.zeros(2 << n, dtype=torch.float64, device="cuda")
    scratch = torch.zeros(int(L.b200k_reduce_scratch_doubles()), dtype=torch.float64, device="cuda")
    dot = torch.zeros(2, dtype=torch.float64, device="cuda")
    scale = float(np.max(np.abs(want)))
    for rep in range(2):  # written, then accumulated
        assert L.b200k_pauli_words_tile(out.data_ptr(), pa, None, n, H, words, nw, rep, None, None, None) == 0
        torch.cuda.synchronize()
        got = out.cpu().numpy().view(np.complex128)
        assert np.max(np.abs(got - (rep + 1) * want)) <= TOL * scale * (rep + 1)
    for bra_ptr, bra in ((pa, va), (pb, vb)):
        assert L.b200k_pauli_words_tile(None, pa, bra_ptr, n, H, words, nw, 0, scratch.data_ptr(), dot.data_ptr(), None) == 0
        torch.cuda.synchronize()
        g = dot.cpu().numpy()
        ref = np.vdot(bra, want)
        assert abs(complex(g[0], g[1]) - ref) <= TOL * max(1.0, abs(ref), scale)
    # argument checks: a word bit outside the tile, too many words, out aliasing in
    bad = (Word * 1)(Word(xin=1 << t, zin=0, zo=0, cr=1.0, ci=0.0, ny=0))
    assert L.b200k_pauli_words_tile(out.data_ptr(), pa, None, n, H, bad, 1, 0, None, None, None) != 0
    assert L.b200k_pauli_words_tile(out.data_ptr(), pa, None, n, H, words, 97, 0, None, None, None) != 0
    assert L.b200k_pauli_words_tile(pa, pa, None, n, H, words, nw, 0, None, None, None) != 0
    da.close()
    db.close()
