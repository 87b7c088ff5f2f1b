"""N1: the PennyLane-side `execute()` walks a tape into the device.  PennyLane is not
installable on this image, so the walker (`catalyst_b200.device.execute_tape`, duck-typed against the
QuantumScript protocol) is driven with stand-in tape classes: on CPU against the oracle with hand-computed
answers, on the GPU (`-m gpu`) against the oracle through capi.B200Device."""
import numpy as np
import pytest

from catalyst_b200.device import execute_tape


class Op:
    def __init__(self, name, params=(), wires=(), **hyper):
        self.name, self.parameters, self.wires, self.hyperparameters = name, list(params), list(wires), hyper


class Adjoint:
    def __init__(self, base):
        self.base, self.name, self.wires, self.parameters = base, f"Adjoint({base.name})", base.wires, base.parameters


class Controlled:
    def __init__(self, base, control_wires, control_values=None):
        self.base, self.name = base, f"C({base.name})"
        self.control_wires, self.control_values = list(control_wires), control_values
        self.wires, self.parameters = list(control_wires) + list(base.wires), base.parameters


class Obs(Op):
    pass


class Prod:
    name = "Prod"

    def __init__(self, *operands):
        self.operands = operands


class SProd:
    name = "SProd"

    def __init__(self, scalar, base):
        self.scalar, self.base = scalar, base


class Sum:
    name = "Sum"

    def __init__(self, coeffs, ops):
        self._c, self._o = coeffs, ops

    def terms(self):
        return self._c, self._o


def _mp(kind, obs=None, wires=None):
    return type(kind, (), {})(), obs, wires


class MP:
    def __init__(self, obs=None, wires=None):
        self.obs, self.wires = obs, wires


ExpectationMP = type("ExpectationMP", (MP,), {})
VarianceMP = type("VarianceMP", (MP,), {})
ProbabilityMP = type("ProbabilityMP", (MP,), {})
StateMP = type("StateMP", (MP,), {})
SampleMP = type("SampleMP", (MP,), {})


class Shots:
    def __init__(self, n):
        self.total_shots = n


class Tape:
    def __init__(self, ops, measurements, shots=None):
        self.operations, self.measurements = ops, measurements
        self.shots = Shots(shots)
        self.wires = sorted({w for o in ops for w in o.wires})


def _tape():
    ops = [Op("Hadamard", [], ["a"]), Op("CNOT", [], ["a", "b"]), Adjoint(Op("RX", [0.4], ["c"])),
           Controlled(Op("RY", [1.1], ["b"]), ["c"], [False]), Op("PauliRot", [0.3], ["a", "c"], pauli_word="XY"),
           Op("QubitUnitary", [np.array([[0, 1], [1, 0]])], ["b"])]
    ham = Sum([0.5, -1.5], [Prod(Obs("PauliZ", wires=["a"]), Obs("PauliX", wires=["c"])), SProd(2.0, Obs("PauliY", wires=["b"]))])
    ms = [ExpectationMP(obs=ham), VarianceMP(obs=Obs("PauliZ", wires=["b"])), ProbabilityMP(wires=["c", "a"]), StateMP()]
    return Tape(ops, ms)


def _direct(dev):
    """The same program issued by hand (what execute_tape must reproduce)."""
    dev.named_op("Hadamard", [], [0]); dev.named_op("CNOT", [], [0, 1]); dev.named_op("RX", [0.4], [2], inverse=True)
    dev.named_op("RY", [1.1], [1], ctrl_wires=[2], ctrl_values=[False])
    dev.named_op("PauliRot", [0.3], [0, 2], pauli_word="XY")
    dev.matrix_op(np.array([[0, 1], [1, 0]], dtype=complex), [1])
    k = dev.obs_hamiltonian([0.5, -1.5], [dev.obs_tensor([dev.obs_named("PauliZ", 0), dev.obs_named("PauliX", 2)]),
                                           dev.obs_hamiltonian([2.0], [dev.obs_named("PauliY", 1)])])
    return dev.expval(k), dev.var(dev.obs_named("PauliZ", 1)), dev.probs([2, 0]), dev.state()


def test_execute_tape_on_the_oracle(oracle_mod):
    got = execute_tape(_tape(), lambda n, shots: oracle_mod.OracleDevice(n, seed=1))
    want = _direct(oracle_mod.OracleDevice(3, seed=1))
    for g, w in zip(got, want):
        assert np.max(np.abs(np.asarray(g) - np.asarray(w))) < 1e-14
    assert abs(np.sum(got[2]) - 1.0) < 1e-12


def test_execute_tape_samples_on_the_oracle(oracle_mod):
    t = Tape([Op("Hadamard", [], [0]), Op("CNOT", [], [0, 1])], [SampleMP(wires=[0, 1])], shots=50)
    smp = execute_tape(t, lambda n, shots: oracle_mod.OracleDevice(n, seed=3))
    assert smp.shape == (50, 2) and np.all(smp[:, 0] == smp[:, 1])


@pytest.mark.gpu
def test_execute_tape_on_the_device(oracle_mod):
    from catalyst_b200 import capi
    got = execute_tape(_tape(), lambda n, shots: capi.B200Device(n, seed=1))
    want = _direct(oracle_mod.OracleDevice(3, seed=1))
    for g, w in zip(got, want):
        assert np.max(np.abs(np.asarray(g) - np.asarray(w))) < 1e-12
