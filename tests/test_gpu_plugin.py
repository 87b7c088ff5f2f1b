"""The drop-in boundary itself: the plugin .so (C++ vtable of QuantumDevice behind
`B200StateVectorFactory`) driven by the REFERENCE's own runtime C API compiled from its sources
(oracle/_ref/librt_capi.so), i.e. the exact call sequence of a @qjit-compiled program, checked
against the CPU oracle and the reference's seeded golden vectors."""
import numpy as np
import pytest

import rt_driver
from dsl import assert_close, apply_ops, build_obs, random_circuit, random_op, random_pauli_obs, random_state
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
        assert_close(d.var(kd), o.var(ko), "plugin: Var")
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
    assert_close(d.state(), o.state(), "plugin: state after measurement")


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


# ---- DataView strides and offsets (runtime/include/DataView.hpp:124-134, row t1 of SURVEY.md 8(a)) --------
def _strided_memref(M, ctype, base, offset, sizes, strides):
    """A memref descriptor over `base` (1-D numpy array of the element type) with an explicit element
    offset and element strides, as MLIR hands non-contiguous views to the runtime (Types.h:84-150)."""
    m = M()
    p = base.ctypes.data_as(rt_driver.C.POINTER(ctype))
    m.allocated, m.aligned, m.offset = p, p, offset
    for i, (sz, st) in enumerate(zip(sizes, strides)):
        m.sizes[i] = sz
        m.strides[i] = st
    return m


def test_runtime_strided_and_offset_memrefs(rt, oracle_mod):
    """State, Probs, PartialProbs, Sample, Gradient and SetState through views with offset != 0 and
    stride != 1: every element must land where the DataView says, and nothing else may be written."""
    C = rt_driver.C
    n, shots = 6, 200
    ops = random_circuit(21, n, 40)
    init = random_state(5, n)
    o = oracle_mod.OracleDevice(n, seed=11)
    d = rt(n, seed=11)                                       # analytic device: State, Probs, PartialProbs
    L = d.L
    # SetState from a strided complex view: element i at base[3 + 2 i]
    src = np.full(3 + 2 * (1 << n), 7.0 + 7.0j, dtype=np.complex128)
    src[3::2][:1 << n] = init
    cplx_t = C.c_double * 2
    mr = _strided_memref(rt_driver._memref(cplx_t, 1), cplx_t, src.view(np.float64).reshape(-1, 2), 3, [1 << n], [2])
    L.qis__SetState(C.byref(mr), C.c_uint64(n), *[d.q[w] for w in range(n)])
    o.set_state(init, list(range(n)))
    apply_ops(o, ops)
    apply_ops(d, ops)
    # State into base[5 + 3 i]
    SENT = -123.0 - 456.0j
    dst = np.full(5 + 3 * (1 << n), SENT, dtype=np.complex128)
    mr = _strided_memref(rt_driver._memref(cplx_t, 1), cplx_t, dst.view(np.float64).reshape(-1, 2), 5, [1 << n], [3])
    L.qis__State(C.byref(mr), C.c_int64(0))
    want = o.state()
    assert np.max(np.abs(dst[5::3][:1 << n] - want)) < TOL
    mask = np.ones(dst.size, dtype=bool)
    mask[5 + 3 * np.arange(1 << n)] = False
    assert np.all(dst[mask] == SENT), "State wrote outside the view"
    # Probs (all wires) into base[1 + 4 i]; PartialProbs on 3 wires into base[2 + 5 i]
    pd = np.full(1 + 4 * (1 << n), -1.0)
    mr = _strided_memref(rt_driver.MR_d1, C.c_double, pd, 1, [1 << n], [4])
    L.qis__Probs(C.byref(mr), C.c_int64(0))
    assert np.max(np.abs(pd[1::4][:1 << n] - o.probs())) < TOL
    assert np.count_nonzero(pd == -1.0) == pd.size - (1 << n)
    w = [4, 0, 2]
    pp = np.full(2 + 5 * 8, -1.0)
    mr = _strided_memref(rt_driver.MR_d1, C.c_double, pp, 2, [8], [5])
    L.qis__Probs(C.byref(mr), C.c_int64(3), *[d.q[x] for x in w])
    assert np.max(np.abs(pp[2::5][:8] - o.probs(w))) < TOL
    d.close()
    # Sample (a device with shots) into a column-major-like 2-D view: element (s, k) at base[7 + 2 s + k (2 shots + 3)]
    o = oracle_mod.OracleDevice(n, seed=11)
    d = rt(n, seed=11, shots=shots)
    L = d.L
    apply_ops(o, ops)
    apply_ops(d, ops)
    st1 = 2 * shots + 3
    sd = np.full(7 + 2 * shots + n * st1, -1.0)
    mr = _strided_memref(rt_driver.MR_d2, C.c_double, sd, 7, [shots, n], [2, st1])
    L.qis__Sample(C.byref(mr), C.c_int64(0))
    got = np.array([[sd[7 + s * 2 + k * st1] for k in range(n)] for s in range(shots)])
    assert np.array_equal(got, o.sample(shots))
    assert np.count_nonzero(sd == -1.0) == sd.size - shots * n


def test_runtime_strided_gradient(rt, oracle_mod):
    C = rt_driver.C
    n = 5
    ops = [("RX", [0.3], [0]), ("RY", [1.1], [1]), ("CNOT", [], [0, 2]), ("IsingZZ", [0.7], [1, 3]), ("CRX", [-0.4], [2, 4]),
           ("RZ", [2.2], [4])]
    obs_list = [("N", "PauliZ", 2), ("T", [("N", "PauliX", 1), ("N", "PauliY", 4)])]
    o = oracle_mod.OracleDevice(n)
    o.start_tape()
    apply_ops(o, ops)
    for ob in obs_list:
        o.expval(build_obs(o, ob))
    o.stop_tape()
    jo = o.gradient()
    d = rt(n)
    d.start_tape()
    apply_ops(d, ops)
    for ob in obs_list:
        d.expval(build_obs(d, ob))
    d.stop_tape()
    P = jo.shape[1]
    rows = [np.full(4 + 3 * P, -9.0) for _ in obs_list]
    mrs = [_strided_memref(rt_driver.MR_d1, C.c_double, r, 4, [P], [3]) for r in rows]
    d.L.qis__Gradient(C.c_int64(len(obs_list)), *[C.byref(m) for m in mrs])
    for k, r in enumerate(rows):
        assert np.max(np.abs(r[4::3][:P] - jo[k])) < TOL * 10
        assert np.count_nonzero(r == -9.0) == r.size - P


# ---- runtime-side wrappers that need nothing new from the device ------------------
def test_runtime_mbqc_measure_in_basis(rt, oracle_mod):
    """__catalyst__mbqc__measure_in_basis (RuntimeCAPI.cpp:1173-1184) is Measure() on the device: same seeded
    outcomes and the same collapsed state as the oracle's Measure."""
    C = rt_driver.C
    n = 5
    ops = random_circuit(3, n, 25)
    o = oracle_mod.OracleDevice(n, seed=37)
    d = rt(n, seed=37)
    apply_ops(o, ops)
    apply_ops(d, ops)
    d.L.mbqc__measure_in_basis.restype = C.POINTER(C.c_bool)
    for wire, post in [(0, -1), (3, -1), (1, 1), (4, -1)]:
        r = d.L.mbqc__measure_in_basis(d.q[wire], C.c_uint32(0), C.c_double(0.25), C.c_int32(post))
        assert bool(r[0]) == o.measure(wire, None if post < 0 else post)
    assert_close(d.state(), o.state(), "plugin: state after measurement")


def test_runtime_print_state(rt, capfd):
    """__catalyst__rt__print_state (RuntimeCAPI.cpp:278-290) = State() of the plugin written to stdout."""
    d = rt(2)
    d.named_op("Hadamard", [], [0])
    d.named_op("CNOT", [], [0, 1])
    d.L.rt__print_state()
    import ctypes
    ctypes.CDLL(None).fflush(None)
    out = capfd.readouterr().out
    assert "State = [" in out
    vals = [ln.strip().rstrip(",") for ln in out.splitlines() if ln.strip().startswith("(")]
    assert len(vals) == 4
    amps = [complex(*map(float, v.strip("()").split(","))) for v in vals]
    assert np.allclose(amps, [2 ** -0.5, 0, 0, 2 ** -0.5])
