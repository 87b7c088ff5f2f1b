"""Parity of the SHARDED device against the CPU oracle, on ONE GPU: the ranks run as
threads of this process joined by the 'local:' rank group ("virtual shards"), which exercises the
same planner, pack/exchange/unpack swap path, rank-bit predicates and cross-rank reductions as the
NCCL build — only the transport differs.  Every rank must return the same, oracle-equal results."""
import numpy as np
import pytest

from dsl import assert_close, apply_ops, build_obs, random_circuit, random_pauli_obs, random_state

pytestmark = pytest.mark.gpu
TOL = 1e-12


@pytest.fixture(scope="module")
def gpu():
    from catalyst_b200 import capi
    return capi


def _oracle(oracle_mod, n, ops, seed=None):
    o = oracle_mod.OracleDevice(n, seed=seed)
    apply_ops(o, ops)
    return o


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("n,depth,seed", [(7, 40, 1), (8, 80, 2), (10, 120, 3), (12, 150, 4)])
def test_state_random_circuits(gpu, oracle_mod, world, n, depth, seed):
    ops = random_circuit(seed * 100 + world, n, depth)
    want = _oracle(oracle_mod, n, ops).state()

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return d.state(), d.stats().exchanges

    res = gpu.run_spmd(world, prog)
    for st, _ in res:
        assert np.max(np.abs(st - want)) < TOL
    assert res[0][1] > 0, "no global<->local swap was exercised"


@pytest.mark.parametrize("world", [2, 4])
def test_swap_chunking_small_stage(gpu, oracle_mod, world):
    """stage_mib=1 at n=18 forces many chunks per swap."""
    n = 18
    ops = random_circuit(77, n, 60, names=["RX", "RY", "Hadamard", "CNOT", "CZ", "RZ", "SWAP"])
    want = _oracle(oracle_mod, n, ops).state()

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return d.state()

    for st in gpu.run_spmd(world, prog, stage_mib=1):
        assert np.max(np.abs(st - want)) < TOL


@pytest.mark.parametrize("world", [2, 4])
def test_expval_var_probs(gpu, oracle_mod, world):
    n = 9
    rng = np.random.default_rng(5)
    ops = random_circuit(11, n, 60)
    obs_list = [random_pauli_obs(rng, n) for _ in range(6)]
    obs_list.append(("Ham", [0.3, -1.2, 0.7], [obs_list[0], obs_list[1], ("N", "Hadamard", 0)]))
    herm = rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4))
    herm = herm + herm.conj().T
    obs_list.append(("H", herm, [0, n - 1]))
    obs_list.append(("T", [("H", herm, [1, 0]), ("N", "PauliZ", 5)]))
    o = _oracle(oracle_mod, n, ops)
    want_e = [o.expval(build_obs(o, ob)) for ob in obs_list]
    want_v = [o.var(build_obs(o, ob)) for ob in obs_list]
    wire_sets = [[0], [n - 1], [0, 1, 2], [3, 0, 8], list(range(n)), [8, 7, 6, 5, 4, 3, 2, 1, 0]]
    want_p = [o.probs(w) for w in wire_sets] + [o.probs()]

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        e = [d.expval(build_obs(d, ob)) for ob in obs_list]
        v = [d.var(build_obs(d, ob)) for ob in obs_list]
        p = [d.probs(w) for w in wire_sets] + [d.probs()]
        return e, v, p, d.state()

    for e, v, p, st in gpu.run_spmd(world, prog, tile_bits=5):
        assert np.max(np.abs(np.array(e) - want_e)) < TOL
        assert_close(np.array(v), want_v, "sharded: Var")
        for a, b in zip(p, want_p):
            assert np.max(np.abs(a - b)) < TOL
        assert np.max(np.abs(st - o.state())) < TOL  # swaps made by the measurements are consistent


@pytest.mark.parametrize("world", [2, 4])
def test_state_prep(gpu, oracle_mod, world):
    n = 7
    full = random_state(3, n)
    sub = random_state(4, 3)
    ops = random_circuit(21, n, 30)

    def prog(d, r):
        out = []
        d.allocate_qubits(n)
        d.set_state(full, list(range(n)))
        apply_ops(d, ops)
        out.append(d.state())
        d.set_basis_state([1, 0, 1], [0, 3, 6])
        apply_ops(d, ops)
        out.append(d.state())
        d.set_state(sub, [6, 0, 2])
        apply_ops(d, ops)
        out.append(d.state())
        return out

    want = []
    o = oracle_mod.OracleDevice(n)
    o.set_state(full, list(range(n))); apply_ops(o, ops); want.append(o.state())
    o.set_basis_state([1, 0, 1], [0, 3, 6]); apply_ops(o, ops); want.append(o.state())
    o.set_state(sub, [6, 0, 2]); apply_ops(o, ops); want.append(o.state())
    for got in gpu.run_spmd(world, prog, tile_bits=4):
        for a, b in zip(got, want):
            assert np.max(np.abs(a - b)) < TOL


@pytest.mark.parametrize("world", [2, 4])
def test_measure_and_sample_bit_exact(gpu, oracle_mod, world):
    n = 8
    ops = random_circuit(31, n, 50)
    o = _oracle(oracle_mod, n, ops, seed=37)
    want_m = [o.measure(w) for w in (0, 5, 7)]
    want_s = o.sample(500)
    want_sp = o.sample(300, [6, 0])
    want_c = o.counts(200, [1, 7, 0])
    want_state = o.state()

    def prog(d, r):
        d.seed(37)
        d.allocate_qubits(n)
        apply_ops(d, ops)
        m = [d.measure(w) for w in (0, 5, 7)]
        return m, d.sample(500), d.sample(300, [6, 0]), d.counts(200, [1, 7, 0]), d.state()

    for m, s, sp, c, st in gpu.run_spmd(world, prog, tile_bits=5):
        assert m == want_m
        assert np.array_equal(s, want_s)
        assert np.array_equal(sp, want_sp)
        assert np.array_equal(c[0], want_c[0]) and np.array_equal(c[1], want_c[1])
        assert np.max(np.abs(st - want_state)) < TOL


@pytest.mark.parametrize("world", [2, 4])
def test_adjoint_gradient(gpu, oracle_mod, world):
    n = 8
    names = ["RX", "RY", "RZ", "PhaseShift", "IsingXX", "IsingYY", "IsingZZ", "IsingXY", "CRX", "CRY",
             "CRZ", "ControlledPhaseShift", "SingleExcitation", "DoubleExcitation", "MultiRZ",
             "CNOT", "Hadamard", "CZ", "SWAP", "Toffoli", "PSWAP"]
    ops = random_circuit(41, n, 45, names=names, allow_ctrl=False)
    rng = np.random.default_rng(9)
    obs_list = [random_pauli_obs(rng, n), ("N", "PauliZ", 0),
                ("Ham", [0.5, -0.25], [random_pauli_obs(rng, n), random_pauli_obs(rng, n)])]

    def run(d):
        d.start_tape()
        apply_ops(d, ops)
        e = [d.expval(build_obs(d, ob)) for ob in obs_list]
        d.stop_tape()
        return e, d.gradient()

    o = oracle_mod.OracleDevice(n)
    want_e, want_g = run(o)

    def prog(d, r):
        d.allocate_qubits(n)
        e, g = run(d)
        return e, g, d.state()

    for e, g, st in gpu.run_spmd(world, prog, tile_bits=5):
        assert np.max(np.abs(np.array(e) - want_e)) < TOL
        assert_close(g, want_g, "sharded: adjoint gradient")
        assert np.max(np.abs(st - o.state())) < TOL


def test_sharded_errors(gpu):
    with pytest.raises(gpu.B200Error):
        gpu.B200Device(world_size=3, rank=0, comm="local:x")
    with pytest.raises(gpu.B200Error):
        gpu.B200Device(world_size=2, rank=0)  # no comm
    with pytest.raises(gpu.B200Error):
        gpu.B200Device(world_size=2, rank=2, comm="local:x")

    def prog(d, r):
        with pytest.raises(gpu.B200Error):
            d.allocate_qubits(3)  # 8 ranks need >= 6 qubits
        d.allocate_qubits(6)
        d.allocate_qubit()                     # dynamic allocation works on a sharded register (a new local bit)
        assert d.num_qubits() == 7
        d.release_qubit(6)
        with pytest.raises(gpu.B200Error):
            d.release_qubit(0)                 # ... but 8 ranks keep at least 6 qubits
        return True

    assert all(gpu.run_spmd(8, prog))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pauli_measure_sharded(gpu, oracle_mod, world):
    """Joint Pauli measurements whose X/Y factors sit on rank bits force swaps; every rank must see the
    same outcome sequence as the oracle and end in the same state."""
    n = 9
    rng = np.random.default_rng(world)
    ops = random_circuit(41 + world, n, 40)
    words = []
    for _ in range(5):
        k = int(rng.integers(1, 5))
        words.append(("".join(rng.choice(list("XYZ"), size=k)), [int(w) for w in rng.permutation(n)[:k]]))
    o = _oracle(oracle_mod, n, ops, seed=37)
    want = [o.pauli_measure(w, ws) for w, ws in words]
    want_state = o.state()

    def prog(d, r):
        d.seed(37)
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return [d.pauli_measure(w, ws) for w, ws in words], d.state()

    for m, st in gpu.run_spmd(world, prog, tile_bits=4, min_row_bits=2):
        assert m == want
        assert_close(st, want_state, "sharded: state after PauliMeasure")


@pytest.mark.parametrize("world", [2, 4])
def test_shots_expval_sharded(gpu, oracle_mod, world):
    n, shots = 8, 3000
    ops = random_circuit(55, n, 40)
    ham = ("Ham", [0.3, -0.8, 1.1], [("T", [("N", "PauliX", 0), ("N", "PauliY", 7)]), ("N", "Hadamard", 1),
                                     ("T", [("N", "PauliZ", 0), ("N", "PauliX", 4)])])
    o = _oracle(oracle_mod, n, ops, seed=37)
    o.set_shots(shots)
    want_e = o.expval(build_obs(o, ham))
    want_p = o.probs([0, 6])

    def prog(d, r):
        d.seed(37)
        d.allocate_qubits(n)
        apply_ops(d, ops)
        d.set_shots(shots)
        return d.expval(build_obs(d, ham)), d.probs([0, 6])

    for e, p in gpu.run_spmd(world, prog, tile_bits=4, min_row_bits=2):
        assert abs(e - want_e) < 1e-12
        assert np.array_equal(p, want_p)


@pytest.mark.parametrize("p2p", [True, False])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_swap_paths_peer_memory_and_staged(gpu, oracle_mod, world, p2p):
    """The global<->local qubit swap has two transports: in-place kernels over the peers' mapped shards
    (default) and pack -> exchange -> unpack through a staging buffer (`p2p_swap=False`, also used for
    gradient work vectors).  Both must give the oracle's state; the statistics say which one ran."""
    n = 11
    ops = random_circuit(500 + world, n, 120)
    want = _oracle(oracle_mod, n, ops).state()

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        st = d.state()
        s = d.stats()
        # release / allocate of the same size keeps the mapping; run once more
        d.release_qubits()
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return st, int(s.exchanges), int(s.p2p_exchanges), d.state()

    for st, ex, p2, st2 in gpu.run_spmd(world, prog, p2p_swap=int(p2p), tile_bits=5, min_row_bits=2):
        assert np.max(np.abs(st - want)) < TOL
        assert np.max(np.abs(st2 - want)) < TOL
        assert ex > 0
        assert (p2 == ex) if p2p else (p2 == 0)


@pytest.mark.parametrize("world", [2, 4])
def test_swap_relabelling_sharded(gpu, oracle_mod, world):
    n = 9
    rng = np.random.default_rng(3)
    ops = []
    for layer in range(5):
        ops += random_circuit(600 + layer, n, 8)
        ops += [("SWAP", [], [int(x) for x in rng.permutation(n)[:2]]) for _ in range(3)]
    o = _oracle(oracle_mod, n, ops, seed=37)
    want_p, want_pp, want_s = o.probs(), o.probs([0, 8, 4]), o.sample(300)
    want = o.state()

    def prog(d, r):
        d.seed(37)
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return d.probs(), d.probs([0, 8, 4]), d.sample(300), d.state(), int(d.stats().relabelled_swaps)

    for p, pp, s, st, rs in gpu.run_spmd(world, prog, tile_bits=4, min_row_bits=2):
        assert rs >= 15
        assert np.max(np.abs(p - want_p)) < TOL and np.max(np.abs(pp - want_pp)) < TOL
        assert np.array_equal(s, want_s)
        assert np.max(np.abs(st - want)) < TOL


@pytest.mark.parametrize("world", [2, 4])
def test_unseeded_measurement_agrees_across_ranks(gpu, world):
    """No PRNG set (the reference default, SetDevicePRNG(nullptr)): every rank must still collapse onto the
    SAME outcome and return the same samples — rank 0's entropy is distributed."""
    n = 8

    def prog(d, r):
        d.allocate_qubits(n)                      # note: no seed
        for w in range(n):
            d.named_op("Hadamard", [], [w])
        outs = [d.measure(w) for w in (0, n - 1, 3)]            # wire 0 sits on a rank bit
        pm = d.pauli_measure("XZ", [1, 2])
        smp = d.sample(64)
        st = d.state()
        return outs, pm, smp, st

    res = gpu.run_spmd(world, prog)
    for outs, pm, smp, st in res[1:]:
        assert outs == res[0][0] and pm == res[0][1]
        assert np.array_equal(smp, res[0][2])
        assert np.array_equal(st, res[0][3])
    st = res[0][3]
    assert abs(np.vdot(st, st).real - 1.0) < 1e-12
    # the measured wires are frozen at their outcomes
    for w, m in zip((0, n - 1, 3), res[0][0]):
        assert np.all(res[0][2][:, w] == float(m))


@pytest.mark.parametrize("stage_mib", [1, 4096])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_swap_push_mode(gpu, oracle_mod, world, stage_mib):
    """Third transport of the qubit swap (default over NCCL process groups): one kernel per chunk pushes this
    rank's sub-blocks into every peer's mapped receive staging and scatters the previous chunk.  stage_mib=1 at
    17 qubits forces several chunks per swap (both staging halves, the trailing unpack-only round); a release /
    allocate at another size re-creates and re-maps the staging; the adjoint gradient swaps several vectors."""
    n = 17 if stage_mib == 1 else 11
    ops = random_circuit(900 + world, n, 80, names=["RX", "RY", "RZ", "Hadamard", "CNOT", "CZ", "IsingXX", "CRY", "T"])
    want = _oracle(oracle_mod, n, ops).state()
    gops = [("RX", [0.3], [0]), ("RY", [0.7], [n - 1]), ("CNOT", [], [0, n - 1]), ("IsingXX", [0.4], [1, n - 2]),
            ("RZ", [1.2], [0]), ("CRX", [0.9], [n - 1, 0])]
    og = oracle_mod.OracleDevice(9)
    og.start_tape()
    apply_ops(og, [(g[0], g[1], [w if w < 4 else 9 - (n - w) for w in g[2]]) for g in gops])
    og.expval(og.obs_tensor([og.obs_named("PauliX", 0), og.obs_named("PauliZ", 8)]))
    og.stop_tape()
    jo = og.gradient()

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        st = d.state()
        s = d.stats()
        d.release_qubits()
        d.allocate_qubits(9)                      # another shard size: staging is re-created and re-mapped
        d.start_tape()
        apply_ops(d, [(g[0], g[1], [w if w < 4 else 9 - (n - w) for w in g[2]]) for g in gops])
        d.expval(d.obs_tensor([d.obs_named("PauliX", 0), d.obs_named("PauliZ", 8)]))
        d.stop_tape()
        return st, int(s.exchanges), int(s.swap_push_exchanges), d.gradient()

    for st, ex, pushed, jd in gpu.run_spmd(world, prog, swap_push=1, p2p_swap=0, stage_mib=stage_mib, tile_bits=6, min_row_bits=2):
        assert np.max(np.abs(st - want)) < TOL
        assert ex > 0 and pushed == ex
        assert np.max(np.abs(jd - jo)) < 1e-12


# ---- Sample / Counts of ALL wires without gathering the marginal (the distributed sampler of SURVEY 8(e)) ----------
def _ghz(n):
    return [("Hadamard", [], [0])] + [("CNOT", [], [w, w + 1]) for w in range(n - 1)]


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("kind", ["random", "ghz", "basis"])
def test_distributed_sampler_all_wires(gpu, oracle_mod, world, kind):
    """dist_sampler=1 REQUIRES the distributed path (an error if it cannot run): every rank keeps |psi|^2 of its own
    slice, chunk sums are gathered, every rank draws every shot through peer memory.  Bit-exact with the oracle's
    (= the unsharded device's) seeded stream, for a spread-out state, for two peaks on different ranks and for a basis
    state; partial-wire Sample between the calls keeps the generator in step."""
    n = 14
    ops = {"random": random_circuit(57, n, 120), "ghz": _ghz(n), "basis": [("PauliX", [], [2]), ("PauliX", [], [n - 1])]}[kind]
    o = _oracle(oracle_mod, n, ops, seed=37)
    want = (o.sample(2000), o.counts(1000), o.sample(300, [3, 0]), o.sample(500), o.state())

    def prog(d, r):
        d.seed(37)
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return d.sample(2000), d.counts(1000), d.sample(300, [3, 0]), d.sample(500), d.state()

    for s, c, sp, s2, st in gpu.run_spmd(world, prog, dist_sampler=1):
        assert np.array_equal(s, want[0])
        assert np.array_equal(c[0], want[1][0]) and np.array_equal(c[1], want[1][1])
        assert np.array_equal(sp, want[2])
        assert np.array_equal(s2, want[3])
        assert np.max(np.abs(st - want[4])) < TOL


def test_distributed_sampler_refused_when_required_but_unavailable(gpu):
    def prog(d, r):
        d.allocate_qubits(6)  # 2^5 amplitudes per rank: less than one chunk
        with pytest.raises(gpu.B200Error, match="dist_sampler=1"):
            d.sample(10)
        return True

    assert all(gpu.run_spmd(2, prog, dist_sampler=1))


@pytest.mark.parametrize("world", [2, 4])
def test_state_and_probs_slices(gpu, oracle_mod, world):
    """gather='slice': State() and Probs() of all wires fill only this rank's slice of the caller's array (no rank ever
    holds 2^n entries); the register stays usable afterwards (its layout was brought to wire order)."""
    n = 12
    nl = n - (world.bit_length() - 1)
    ops = random_circuit(61, n, 100)
    o = _oracle(oracle_mod, n, ops)
    want_state, want_probs = o.state(), o.probs()
    apply_ops(o, [("Hadamard", [], [n - 1]), ("CNOT", [], [0, n - 1])])
    want_e = o.expval(o.obs_named("PauliZ", 0))
    want_after = o.state()

    def prog(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        p = np.full(1 << n, np.nan)
        d.probs(out=p)
        st = np.full(1 << n, np.nan + 0j, dtype=np.complex128)
        d.state(out=st)
        apply_ops(d, [("Hadamard", [], [n - 1]), ("CNOT", [], [0, n - 1])])
        e = d.expval(d.obs_named("PauliZ", 0))
        st2 = np.full(1 << n, np.nan + 0j, dtype=np.complex128)
        d.state(out=st2)
        return st, p, e, st2

    for r, (st, p, e, st2) in enumerate(gpu.run_spmd(world, prog, gather="slice")):
        lo, hi = r << nl, (r + 1) << nl
        assert np.max(np.abs(st[lo:hi] - want_state[lo:hi])) < TOL
        assert np.max(np.abs(p[lo:hi] - want_probs[lo:hi])) < TOL
        assert np.max(np.abs(st2[lo:hi] - want_after[lo:hi])) < TOL
        mask = np.ones(1 << n, dtype=bool)
        mask[lo:hi] = False
        assert np.all(np.isnan(st[mask].real)) and np.all(np.isnan(p[mask])), "entries outside the slice were written"
        assert abs(e - want_e) < TOL
