"""Helpers to run the tests' mini-DSL (see tests/golden/reference_cases.py) on any object that
implements the device method set (oracle.OracleDevice or catalyst_b200.capi.B200Device)."""
import numpy as np


# ---- the parity bar of BASELINE.json's north_star: 1e-12 RELATIVE --------------------------------------
ACHIEVED = {}   # label -> largest relative error seen (printed at the end of the session by conftest.py)


def assert_close(got, want, label, tol=1e-12):
    """max |got - want| <= tol * max(1, max |want|); records the achieved error under `label`."""
    got, want = np.asarray(got), np.asarray(want)
    scale = max(1.0, float(np.max(np.abs(want)))) if want.size else 1.0
    err = float(np.max(np.abs(got - want))) / scale if want.size else 0.0
    ACHIEVED[label] = max(ACHIEVED.get(label, 0.0), err)
    assert err <= tol, f"{label}: relative error {err:.3e} exceeds {tol:g}"


def apply_ops(dev, ops):
    for op in ops:
        name, params, wires = op[0], op[1], op[2]
        kw = op[3] if len(op) > 3 else {}
        if name == "QubitUnitary":
            dev.matrix_op(kw["matrix"], wires, inverse=kw.get("inverse", False),
                          ctrl_wires=kw.get("ctrl_wires", ()), ctrl_values=kw.get("ctrl_values", ()))
        else:
            dev.named_op(name, params, wires, inverse=kw.get("inverse", False),
                         ctrl_wires=kw.get("ctrl_wires", ()), ctrl_values=kw.get("ctrl_values", ()),
                         pauli_word=kw.get("pauli_word"))


def build_obs(dev, obs):
    kind = obs[0]
    if kind == "N":
        return dev.obs_named(obs[1], obs[2])
    if kind == "H":
        return dev.obs_hermitian(np.asarray(obs[1], dtype=np.complex128), obs[2])
    if kind == "T":
        return dev.obs_tensor([build_obs(dev, o) for o in obs[1]])
    if kind == "Ham":
        return dev.obs_hamiltonian(list(obs[1]), [build_obs(dev, o) for o in obs[2]])
    raise ValueError(kind)


# ---- random circuit generators (seeded; shared by CPU and GPU tests) -----------------------------
GATES_1Q_0P = ["PauliX", "PauliY", "PauliZ", "Hadamard", "S", "T"]
GATES_1Q_1P = ["PhaseShift", "RX", "RY", "RZ"]
GATES_2Q_0P = ["CNOT", "CY", "CZ", "SWAP", "ISWAP"]
GATES_2Q_1P = ["PSWAP", "IsingXX", "IsingYY", "IsingXY", "IsingZZ", "ControlledPhaseShift",
               "CRX", "CRY", "CRZ", "SingleExcitation"]
GATES_3Q_0P = ["CSWAP", "Toffoli"]
ALL_FIXED = (GATES_1Q_0P + GATES_1Q_1P + ["Rot"] + GATES_2Q_0P + GATES_2Q_1P + ["CRot"]
             + GATES_3Q_0P + ["DoubleExcitation"])
ARITY = {**{g: (1, 0) for g in GATES_1Q_0P}, **{g: (1, 1) for g in GATES_1Q_1P}, "Rot": (1, 3),
         **{g: (2, 0) for g in GATES_2Q_0P}, **{g: (2, 1) for g in GATES_2Q_1P}, "CRot": (2, 3),
         **{g: (3, 0) for g in GATES_3Q_0P}, "DoubleExcitation": (4, 1)}
# gates the reference TOML marks "controllable" (runtime/lib/backend/null_qubit/null_qubit.toml:25-63)
CONTROLLABLE = ["PauliX", "PauliY", "PauliZ", "Hadamard", "S", "T", "PhaseShift", "RX", "RY", "RZ",
                "Rot", "SWAP", "IsingXX", "IsingYY", "IsingXY", "IsingZZ", "SingleExcitation",
                "DoubleExcitation", "MultiRZ", "GlobalPhase", "PauliRot", "PCPhase"]


def random_unitary(rng, k):
    d = 1 << k
    a = rng.normal(size=(d, d)) + 1j * rng.normal(size=(d, d))
    q, r = np.linalg.qr(a)
    return q * (np.diag(r) / np.abs(np.diag(r)))


def random_op(rng, n, allow_ctrl=True, names=None):
    """One random operation on n wires drawn from the full runtime gate set."""
    pool = list(names) if names else (ALL_FIXED + ["MultiRZ", "GlobalPhase", "PCPhase", "PauliRot",
                                                   "QubitUnitary", "Identity"])
    while True:
        name = pool[rng.integers(len(pool))]
        if name in ARITY:
            nw, npar = ARITY[name]
        elif name in ("MultiRZ", "PauliRot"):
            nw, npar = int(rng.integers(1, min(n, 5) + 1)), 1
        elif name == "PCPhase":
            nw, npar = int(rng.integers(1, min(n, 3) + 1)), 2
        elif name == "GlobalPhase":
            nw, npar = 0, 1
        elif name == "Identity":
            nw, npar = int(rng.integers(1, min(n, 2) + 1)), 0
        elif name == "QubitUnitary":
            nw, npar = int(rng.integers(1, min(n, 3) + 1)), 0
        if nw <= n:
            break
    perm = rng.permutation(n)
    wires = [int(w) for w in perm[:nw]]
    rest = [int(w) for w in perm[nw:]]
    params = [float(x) for x in rng.uniform(-2 * np.pi, 2 * np.pi, size=npar)]
    kw = {"inverse": bool(rng.integers(2))}
    if name == "PCPhase":
        params[1] = float(rng.integers(0, (1 << nw) + 1))
    if name == "PauliRot":
        kw["pauli_word"] = "".join("IXYZ"[i] for i in rng.integers(0, 4, size=nw))
    if name == "QubitUnitary":
        kw["matrix"] = random_unitary(rng, nw)
    can_ctrl = allow_ctrl and (name in CONTROLLABLE or name == "QubitUnitary")
    if can_ctrl and rest and rng.random() < 0.4:
        nc = int(rng.integers(1, min(len(rest), 3) + 1))
        kw["ctrl_wires"] = rest[:nc]
        kw["ctrl_values"] = [bool(b) for b in rng.integers(0, 2, size=nc)]
    return (name, params, wires, kw)


def random_circuit(seed, n, depth, **kw):
    rng = np.random.default_rng(seed)
    return [random_op(rng, n, **kw) for _ in range(depth)]


def random_state(seed, n):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    return v / np.linalg.norm(v)


def random_pauli_obs(rng, n, p_identity=0.5):
    """Tensor of single-wire named Paulis (at least one non-identity)."""
    while True:
        word = rng.choice(["I", "X", "Y", "Z"], size=n, p=[p_identity] + [(1 - p_identity) / 3] * 3)
        if any(c != "I" for c in word):
            break
    names = {"X": "PauliX", "Y": "PauliY", "Z": "PauliZ"}
    terms = [("N", names[c], int(w)) for w, c in enumerate(word) if c != "I"]
    return terms[0] if len(terms) == 1 else ("T", terms)


# ---- dynamic qubits: the runtime's id bookkeeping over a device (QubitManager.hpp:110-126) -----------
class DynamicRegister:
    """Runs a program with `qml.allocate` / `deallocate` on a device object the way the runtime does:
    qubit ids are never reused, a new qubit takes device wire n, a release removes the wire and every
    later device wire moves down by one (reference: runtime/lib/backend/common/QubitManager.hpp:110-126;
    callers runtime/lib/capi/RuntimeCAPI.cpp:305-372)."""

    def __init__(self, dev, n_static):
        self.dev = dev
        self.ids = list(range(n_static))        # position in this list = device wire
        self.next_id = n_static

    def wires(self, ids):
        return [self.ids.index(i) for i in ids]

    def allocate(self, k=1):
        out = []
        for _ in range(k):
            self.dev.allocate_qubit()
            self.ids.append(self.next_id)
            out.append(self.next_id)
            self.next_id += 1
        return out

    def release(self, ids):
        for i in ids:
            self.dev.release_qubit(self.ids.index(i))
            self.ids.remove(i)

    def op(self, name, params, ids, ctrl=(), ctrl_values=(), inverse=False):
        self.dev.named_op(name, params, self.wires(ids), inverse=inverse, ctrl_wires=self.wires(ctrl),
                          ctrl_values=list(ctrl_values))

    def measure(self, i, postselect=None, reset=False):
        m = self.dev.measure(self.ids.index(i), postselect)
        if reset and m:                          # catalyst.measure(reset=True): X on outcome 1
            self.op("PauliX", [], [i])
        return m

    def probs(self, ids):
        return self.dev.probs(self.wires(ids))
