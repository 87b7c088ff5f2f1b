"""Internal-consistency checks of the CPU oracle that do not depend on the CUDA device:
operator identities between independently-coded gate paths, unitarity / inverse round trips, and
the adjoint Jacobian against central finite differences (gradients have no in-tree constants in
the reference — SURVEY.md 8(c) "where parity is unpinned")."""
import numpy as np
import pytest

from dsl import ARITY, apply_ops, build_obs, random_circuit, random_op, random_state

N = 5


def run(om, n, ops, init=None):
    dev = om.OracleDevice(n)
    if init is not None:
        dev.set_state(init, list(range(n)))
    apply_ops(dev, ops)
    return dev.state()


@pytest.mark.parametrize("seed", range(30))
def test_random_circuit_unitary_and_inverse(oracle_mod, seed):
    ops = random_circuit(seed, N, 25)
    init = random_state(seed + 1000, N)
    out = run(oracle_mod, N, ops, init)
    assert abs(np.linalg.norm(out) - 1.0) < 1e-12
    inv = []
    for op in reversed(ops):
        kw = dict(op[3])
        kw["inverse"] = not kw.get("inverse", False)
        inv.append((op[0], op[1], op[2], kw))
    back = run(oracle_mod, N, ops + inv, init)
    assert np.allclose(back, init, atol=1e-12)


EQUIV = [
    # (lhs ops, rhs ops) acting on 4 wires
    ([("CNOT", [], [0, 2])], [("PauliX", [], [2], {"ctrl_wires": [0], "ctrl_values": [True]})]),
    ([("CY", [], [3, 1])], [("PauliY", [], [1], {"ctrl_wires": [3], "ctrl_values": [True]})]),
    ([("CZ", [], [1, 2])], [("PauliZ", [], [2], {"ctrl_wires": [1], "ctrl_values": [True]})]),
    ([("Toffoli", [], [0, 1, 3])],
     [("PauliX", [], [3], {"ctrl_wires": [0, 1], "ctrl_values": [True, True]})]),
    ([("CSWAP", [], [2, 0, 3])], [("SWAP", [], [0, 3], {"ctrl_wires": [2], "ctrl_values": [True]})]),
    ([("CRX", [0.7], [0, 1])], [("RX", [0.7], [1], {"ctrl_wires": [0], "ctrl_values": [True]})]),
    ([("CRY", [0.7], [2, 1])], [("RY", [0.7], [1], {"ctrl_wires": [2], "ctrl_values": [True]})]),
    ([("CRZ", [0.7], [3, 0])], [("RZ", [0.7], [0], {"ctrl_wires": [3], "ctrl_values": [True]})]),
    ([("CRot", [0.3, 0.7, -1.1], [3, 0])],
     [("Rot", [0.3, 0.7, -1.1], [0], {"ctrl_wires": [3], "ctrl_values": [True]})]),
    ([("ControlledPhaseShift", [0.9], [1, 3])],
     [("PhaseShift", [0.9], [3], {"ctrl_wires": [1], "ctrl_values": [True]})]),
    ([("Rot", [0.3, 0.7, -1.1], [2])], [("RZ", [0.3], [2]), ("RY", [0.7], [2]), ("RZ", [-1.1], [2])]),
    ([("IsingXX", [1.3], [0, 3])], [("PauliRot", [1.3], [0, 3], {"pauli_word": "XX"})]),
    ([("IsingYY", [1.3], [2, 1])], [("PauliRot", [1.3], [2, 1], {"pauli_word": "YY"})]),
    ([("IsingZZ", [1.3], [2, 3])], [("MultiRZ", [1.3], [2, 3])]),
    ([("MultiRZ", [0.4], [0, 2, 3])], [("PauliRot", [0.4], [0, 2, 3], {"pauli_word": "ZZZ"})]),
    ([("RX", [0.4], [1])], [("PauliRot", [0.4], [1], {"pauli_word": "X"})]),
    ([("RY", [0.4], [1])], [("PauliRot", [0.4], [1], {"pauli_word": "Y"})]),
    ([("RZ", [0.4], [1])], [("MultiRZ", [0.4], [1])]),
    ([("S", [], [1])], [("PhaseShift", [np.pi / 2], [1])]),
    ([("T", [], [1])], [("PhaseShift", [np.pi / 4], [1])]),
    ([("PauliZ", [], [1])], [("PhaseShift", [np.pi], [1])]),
    ([("SWAP", [], [0, 2])], [("CNOT", [], [0, 2]), ("CNOT", [], [2, 0]), ("CNOT", [], [0, 2])]),
    ([("ISWAP", [], [0, 2])], [("PSWAP", [np.pi / 2], [0, 2])]),
    ([("SWAP", [], [1, 2])], [("PSWAP", [0.0], [1, 2])]),
    # IsingXY = exp(i theta/4 (XX+YY)) ; XX and YY commute
    ([("IsingXY", [0.8], [1, 3])],
     [("PauliRot", [-0.4], [1, 3], {"pauli_word": "XX"}), ("PauliRot", [-0.4], [1, 3], {"pauli_word": "YY"})]),
    # SingleExcitation = exp(-i theta/4 (YX - XY))   (first letter acts on first wire)
    ([("SingleExcitation", [0.8], [0, 1])],
     [("PauliRot", [0.4], [0, 1], {"pauli_word": "YX"}), ("PauliRot", [-0.4], [0, 1], {"pauli_word": "XY"})]),
    # GlobalPhase(phi) = exp(-i phi); controlled GlobalPhase = phase on the control subspace
    ([("GlobalPhase", [0.6], [], {"ctrl_wires": [2], "ctrl_values": [True]})], [("PhaseShift", [-0.6], [2])]),
    # PCPhase(theta, dim=2^(k-1)) on k wires = RZ(-2 theta) on the first wire
    ([("PCPhase", [0.35, 2.0], [1, 3])], [("RZ", [-0.7], [1])]),
    # QubitUnitary of a known matrix
    ([("QubitUnitary", [], [2], {"matrix": np.array([[0, 1], [1, 0]])})], [("PauliX", [], [2])]),
    ([("QubitUnitary", [], [0, 3], {"matrix": np.array(
        [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 0, 1], [0, 0, 1, 0]])})], [("CNOT", [], [0, 3])]),
]


@pytest.mark.parametrize("i", range(len(EQUIV)))
def test_operator_identities(oracle_mod, i):
    lhs, rhs = EQUIV[i]
    init = random_state(i, 4)
    a = run(oracle_mod, 4, lhs, init)
    b = run(oracle_mod, 4, rhs, init)
    assert np.allclose(a, b, atol=1e-13), (lhs, rhs)


def test_double_excitation_action(oracle_mod):
    th = 0.73
    c, s = np.cos(th / 2), np.sin(th / 2)
    for src, (d1, a1, d2, a2) in {0b0011: (0b0011, c, 0b1100, s), 0b1100: (0b1100, c, 0b0011, -s)}.items():
        dev = oracle_mod.OracleDevice(4)
        dev.set_basis_state([(src >> 3) & 1, (src >> 2) & 1, (src >> 1) & 1, src & 1], [0, 1, 2, 3])
        dev.named_op("DoubleExcitation", [th], [0, 1, 2, 3])
        st = dev.state()
        exp = np.zeros(16, complex)
        exp[d1], exp[d2] = a1, a2
        assert np.allclose(st, exp)


def test_wire0_is_msb(oracle_mod):
    dev = oracle_mod.OracleDevice(3)
    dev.named_op("PauliX", [], [0])
    assert np.argmax(np.abs(dev.state())) == 4
    assert np.allclose(dev.probs([2, 0]), [0, 1, 0, 0])  # index built from wires in given order


def test_set_state_partial(oracle_mod):
    dev = oracle_mod.OracleDevice(3)
    sub = random_state(3, 2)
    dev.set_state(sub, [2, 0])  # wire 2 is the MSB of the sub-state index
    full = dev.state().reshape(2, 2, 2)  # [w0, w1, w2]
    for a in range(2):
        for b in range(2):
            assert np.isclose(full[b, 0, a], sub[a * 2 + b])
    assert np.allclose(full[:, 1, :], 0)


DIFF_GATES = ["RX", "RY", "RZ", "PhaseShift", "IsingXX", "IsingYY", "IsingXY", "IsingZZ",
              "ControlledPhaseShift", "CRX", "CRY", "CRZ", "SingleExcitation", "DoubleExcitation",
              "MultiRZ", "PauliRot", "GlobalPhase", "PSWAP"]


def _grad_program(om, n, ops, obs_list, thetas=None):
    dev = om.OracleDevice(n)
    dev.set_state(random_state(99, n), list(range(n)))
    dev.start_tape()
    k = 0
    for op in ops:
        name, params, wires, kw = op
        if params and thetas is not None:
            params = [thetas[k]]
            k += 1
        apply_ops(dev, [(name, params, wires, kw)])
    vals = [dev.expval(build_obs(dev, o)) for o in obs_list]
    dev.stop_tape()
    return dev, np.array(vals)


@pytest.mark.parametrize("seed", range(12))
def test_adjoint_gradient_vs_finite_differences(oracle_mod, seed):
    rng = np.random.default_rng(seed)
    n = 5
    ops = []
    for _ in range(14):
        if rng.random() < 0.7:
            name = DIFF_GATES[rng.integers(len(DIFF_GATES))]
            op = random_op(rng, n, names=[name])
        else:
            op = random_op(rng, n, names=["Hadamard", "CNOT", "S", "QubitUnitary", "Toffoli", "SWAP"])
        ops.append(op)
    obs_list = [("N", "PauliZ", 0), ("T", [("N", "PauliX", 1), ("N", "PauliY", 3)]),
                ("Ham", [0.3, -0.8], [("N", "PauliZ", 2), ("H", [[1, 2], [2, -1]], [4])])]
    thetas = np.array([op[1][0] for op in ops if op[1]])
    dev, _ = _grad_program(oracle_mod, n, ops, obs_list, thetas)
    jac = dev.gradient()
    assert jac.shape == (len(obs_list), len(thetas))
    h = 1e-6
    for j in range(len(thetas)):
        tp, tm = thetas.copy(), thetas.copy()
        tp[j] += h
        tm[j] -= h
        _, vp = _grad_program(oracle_mod, n, ops, obs_list, tp)
        _, vm = _grad_program(oracle_mod, n, ops, obs_list, tm)
        fd = (vp - vm) / (2 * h)
        assert np.allclose(jac[:, j], fd, atol=2e-8), (j, ops[j], jac[:, j], fd)
    # trainable subset
    if len(thetas) >= 3:
        sub = [0, len(thetas) - 1]
        assert np.allclose(dev.gradient(sub), jac[:, sub])


def test_gradient_rejects_var_and_multiparam(oracle_mod):
    dev = oracle_mod.OracleDevice(2)
    dev.start_tape()
    dev.named_op("RX", [0.3], [0])
    dev.var(dev.obs_named("PauliZ", 0))
    dev.stop_tape()
    with pytest.raises(RuntimeError):
        dev.gradient()
    dev = oracle_mod.OracleDevice(2)
    dev.start_tape()
    dev.named_op("Rot", [0.3, 0.1, 0.2], [0])
    dev.expval(dev.obs_named("PauliZ", 0))
    dev.stop_tape()
    with pytest.raises(RuntimeError):
        dev.gradient()


# ---- PauliMeasure (QuantumDevice.hpp:531): joint Pauli-product measurement -------------------------
def test_pauli_measure_stabilisers_are_deterministic(oracle_mod):
    # Bell state |00> + |11>: stabilised by +XX, +ZZ, -YY; GHZ: +XXX, +ZZI, +IZZ
    o = oracle_mod.OracleDevice(2, seed=11)
    o.named_op("Hadamard", [], [0]); o.named_op("CNOT", [], [0, 1])
    before = o.state()
    assert [o.pauli_measure("ZZ", [0, 1]), o.pauli_measure("XX", [0, 1]), o.pauli_measure("YY", [0, 1])] == [False, False, True]
    assert np.allclose(o.state(), before)
    o = oracle_mod.OracleDevice(3, seed=11)
    o.named_op("Hadamard", [], [0]); o.named_op("CNOT", [], [0, 1]); o.named_op("CNOT", [], [1, 2])
    assert [o.pauli_measure("XXX", [0, 1, 2]), o.pauli_measure("ZZ", [0, 1]), o.pauli_measure("ZZ", [2, 1]),
            o.pauli_measure("III", [0, 1, 2])] == [False] * 4


@pytest.mark.parametrize("seed", range(6))
def test_pauli_measure_z_equals_measure(oracle_mod, seed):
    from dsl import apply_ops, random_circuit
    n = 5
    ops = random_circuit(seed, n, 25)
    a, b = oracle_mod.OracleDevice(n, seed=seed), oracle_mod.OracleDevice(n, seed=seed)
    apply_ops(a, ops); apply_ops(b, ops)
    for w in (2, 0, 4):
        assert a.pauli_measure("Z", [w]) == b.measure(w)
        assert np.allclose(a.state(), b.state(), atol=1e-14)


@pytest.mark.parametrize("seed", range(8))
def test_pauli_measure_projects_onto_eigenspace(oracle_mod, seed):
    from dsl import apply_ops, build_obs, random_circuit
    rng = np.random.default_rng(seed)
    n = 6
    o = oracle_mod.OracleDevice(n, seed=seed)
    apply_ops(o, random_circuit(seed + 50, n, 30))
    k = int(rng.integers(1, 5))
    wires = [int(w) for w in rng.permutation(n)[:k]]
    word = "".join(rng.choice(list("XYZ"), size=k))
    names = {"X": "PauliX", "Y": "PauliY", "Z": "PauliZ"}
    obs = ("T", [("N", names[c], w) for c, w in zip(word, wires)])
    ev_before = o.expval(build_obs(o, obs))
    r = o.pauli_measure(word, wires)
    assert abs(np.linalg.norm(o.state()) - 1) < 1e-12
    assert abs(o.expval(build_obs(o, obs)) - (-1.0 if r else 1.0)) < 1e-12
    st = o.state()
    assert o.pauli_measure(word, wires) == r and np.allclose(o.state(), st)   # idempotent
    # postselecting the other branch of a fresh copy has probability (1 -+ <P>)/2 > 0 unless deterministic
    if abs(abs(ev_before) - 1) > 1e-9:
        o2 = oracle_mod.OracleDevice(n, seed=seed)
        apply_ops(o2, random_circuit(seed + 50, n, 30))
        assert o2.pauli_measure(word, wires, postselect=int(not r)) == (not r)


def test_pauli_measure_errors(oracle_mod):
    o = oracle_mod.OracleDevice(2)
    with pytest.raises(RuntimeError):
        o.pauli_measure("XX", [0])
    with pytest.raises(RuntimeError):
        o.pauli_measure("Q", [0])
    with pytest.raises(RuntimeError):
        o.pauli_measure("Z", [0], postselect=1)   # |0>: probability of -1 is 0


# ---- finite shots: sampled Expval / Var / Probs -----------------------------------
def _shots_setup(mod, n, seed, shots, dev_seed=7):
    from dsl import apply_ops, random_circuit
    o = mod.OracleDevice(n, seed=dev_seed)
    apply_ops(o, random_circuit(seed, n, 30))
    o.set_shots(shots)
    return o


@pytest.mark.parametrize("seed", range(4))
def test_shots_expval_converges_and_is_deterministic(oracle_mod, seed):
    from dsl import build_obs
    n, shots = 5, 200000
    obs = ("Ham", [0.7, -1.3, 0.4, 2.0],
           [("T", [("N", "PauliX", 0), ("N", "PauliY", 3)]), ("N", "PauliZ", 1), ("N", "Hadamard", 4),
            ("T", [("N", "Identity", 2), ("N", "PauliZ", 2), ("N", "PauliX", 1)])])
    a = _shots_setup(oracle_mod, n, seed, shots)
    b = _shots_setup(oracle_mod, n, seed, shots)
    ea, eb = a.expval(build_obs(a, obs)), b.expval(build_obs(b, obs))
    assert ea == eb                                        # same seed, same stream, same estimate
    a.set_shots(0)
    exact = a.expval(build_obs(a, obs))
    assert abs(ea - exact) < 5 * 4.4 / np.sqrt(shots)      # 5 sigma of the worst-case variance
    # the state is untouched by a sampled expectation value
    c = _shots_setup(oracle_mod, n, seed, 0)
    assert np.allclose(a.state(), c.state(), atol=1e-14)


def test_shots_var_and_probs(oracle_mod):
    from dsl import build_obs
    n, shots = 4, 100000
    o = _shots_setup(oracle_mod, n, 3, shots)
    obs = ("T", [("N", "PauliX", 0), ("N", "PauliZ", 2)])
    v = o.var(build_obs(o, obs))
    p = o.probs([2, 0])
    assert abs(p.sum() - 1) < 1e-12 and np.all(p * shots == np.round(p * shots))
    o.set_shots(0)
    assert abs(v - o.var(build_obs(o, obs))) < 0.03
    assert np.max(np.abs(p - o.probs([2, 0]))) < 0.01
    o.set_shots(10)
    with pytest.raises(RuntimeError):
        o.var(build_obs(o, ("Ham", [1.0, 1.0], [("N", "PauliX", 0), ("N", "PauliZ", 1)])))
    # Hermitian observables are sampled in their eigenbasis: the identity has the single eigenvalue 1
    assert o.expval(build_obs(o, ("H", np.eye(2), [0]))) == 1.0


def test_shots_hermitian_expval_and_var_converge(oracle_mod):
    """Finite-shot Hermitian observables in the oracle: sampled in the eigenbasis of the matrix, the estimates
    converge to the analytic Expval / Var (5 sigma) and are deterministic for a fixed device seed."""
    n, shots = 4, 200000
    rng = np.random.default_rng(0)
    H = rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4))
    H = H + H.conj().T

    def make():
        o = oracle_mod.OracleDevice(n, seed=5)
        for w in range(n):
            o.named_op("RY", [0.3 + w], [w])
        o.named_op("CNOT", [], [0, 2])
        o.named_op("CNOT", [], [1, 3])
        return o, o.obs_tensor([o.obs_hermitian(H, [2, 0]), o.obs_named("PauliX", 3)])
    a, ka = make()
    exact_e, exact_v = a.expval(ka), a.var(ka)
    a.set_shots(shots)
    ea, va = a.expval(ka), a.var(ka)
    b, kb = make()
    b.set_shots(shots)
    assert (b.expval(kb), b.var(kb)) == (ea, va)
    assert abs(ea - exact_e) < 5 * np.sqrt(exact_v / shots)
    assert abs(va - exact_v) < 0.05 * exact_v
    a.set_shots(0)
    assert abs(a.expval(ka) - exact_e) < 1e-13            # the state was restored
