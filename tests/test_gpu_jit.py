"""Run-time specialised tile passes (catalyst_b200/csrc/jit.hpp, simulator_jit.cpp) against the CPU
oracle: every pass of these circuits is compiled by NVRTC into a straight-line sm_100a kernel
(`jit='always'`), and the statistics prove that the specialised kernels, not the interpreted one,
produced the result.  Same tolerance as the interpreted path: 1e-12 on amplitudes."""
import numpy as np
import pytest

from dsl import assert_close, apply_ops, build_obs, random_circuit, random_state

pytestmark = pytest.mark.gpu
TOL = 1e-12

NO_DENSE = ["PauliX", "PauliY", "PauliZ", "Hadamard", "S", "T", "PhaseShift", "RX", "RY", "RZ", "Rot", "CNOT",
            "CY", "CZ", "SWAP", "ISWAP", "PSWAP", "IsingXX", "IsingYY", "IsingXY", "IsingZZ", "SingleExcitation",
            "DoubleExcitation", "ControlledPhaseShift", "CRX", "CRY", "CRZ", "CRot", "CSWAP", "Toffoli", "MultiRZ",
            "GlobalPhase", "PCPhase", "PauliRot"]


FIT2 = [g for g in NO_DENSE if g not in ("DoubleExcitation", "PauliRot")]


@pytest.fixture(scope="module")
def gpu():
    from catalyst_b200 import capi
    return capi


def _run(gpu, oracle_mod, n, ops, init=None, **kw):
    o = oracle_mod.OracleDevice(n)
    d = gpu.B200Device(n, jit="always", **kw)
    if init is not None:
        o.set_state(init, list(range(n)))
        d.set_state(init, list(range(n)))
    apply_ops(o, ops)
    apply_ops(d, ops)
    a, b = o.state(), d.state()
    return d, float(np.max(np.abs(a - b)))


CFG = [dict(), dict(jit_regbits=3), dict(jit_regbits=5), dict(jit_dbuf=2), dict(tile_bits=6, min_row_bits=2),
       dict(tile_bits=12, jit_dbuf=2), dict(tile_bits=9, jit_regbits=2, min_row_bits=3, jit_dbuf=2)]


@pytest.mark.parametrize("cfg", range(len(CFG)))
@pytest.mark.parametrize("n,seed", [(3, 1), (6, 2), (9, 3), (12, 4), (14, 5), (16, 6)])
def test_jit_random_circuit_state(gpu, oracle_mod, n, seed, cfg):
    # gates with at most two non-diagonal bits fit the register rounds of every configuration:
    # every pass must run on a specialised kernel
    ops = random_circuit(seed, n, 50, names=FIT2)
    d, err = _run(gpu, oracle_mod, n, ops, init=random_state(seed + 900, n), **CFG[cfg])
    s = d.stats()
    assert err < TOL
    assert s.jit_passes == s.tile_passes > 0, (s.jit_passes, s.tile_passes)


@pytest.mark.parametrize("cfg", range(len(CFG)))
@pytest.mark.parametrize("n,seed", [(6, 2), (9, 3), (12, 4), (15, 5)])
def test_jit_random_circuit_all_gates(gpu, oracle_mod, n, seed, cfg):
    # a pass holding an op with more non-diagonal bits than register bits (DoubleExcitation, long
    # Pauli words) legitimately stays on the interpreter: parity must hold either way
    ops = random_circuit(seed, n, 60, names=NO_DENSE)
    d, err = _run(gpu, oracle_mod, n, ops, init=random_state(seed + 900, n), **CFG[cfg])
    assert err < TOL


def test_jit_mixed_with_dense(gpu, oracle_mod):
    # MatrixOperation (k >= 2) passes stay on the generic kernel, everything else is specialised
    n = 13
    ops = random_circuit(11, n, 80)
    d, err = _run(gpu, oracle_mod, n, ops, init=random_state(3, n))
    assert err < TOL
    assert d.stats().jit_passes > 0


@pytest.mark.parametrize("cfg", [dict(), dict(jit_regbits=5, tile_bits=12), dict(jit_dbuf=2, tile_bits=7, min_row_bits=3)])
@pytest.mark.parametrize("n", [5, 10, 14])
def test_jit_matrix_operation_in_registers(gpu, oracle_mod, n, cfg):
    """MatrixOperation on k <= jit_regbits wires (with and without controls) is applied as a dense
    2^k x 2^k block on the thread's registers inside the specialised pass."""
    from dsl import random_unitary
    rng = np.random.default_rng(n)
    kmax = min(4, n - 1)
    ops = [("Hadamard", [], [w]) for w in range(n)]
    for i in range(24):
        k = int(rng.integers(2, kmax + 1))
        perm = [int(w) for w in rng.permutation(n)]
        kw = {"matrix": random_unitary(rng, k), "inverse": bool(rng.integers(2))}
        if i % 3 == 0 and n > k + 1:
            kw["ctrl_wires"] = perm[k:k + 2][: n - k]
            kw["ctrl_values"] = [bool(b) for b in rng.integers(0, 2, size=len(kw["ctrl_wires"]))]
        ops.append(("QubitUnitary", [], perm[:k], kw))
        ops.append(("RX", [float(rng.uniform(-3, 3))], [perm[-1]]))
    d, err = _run(gpu, oracle_mod, n, ops, **cfg)
    s = d.stats()
    assert err < TOL
    if cfg.get("tile_bits", 11) - 1 >= 4 and n - 1 >= 4:
        assert s.jit_passes == s.tile_passes > 0, (s.jit_passes, s.tile_passes)


@pytest.mark.parametrize("n", [18, 21])
def test_jit_qft_random_layers(gpu, oracle_mod, n):
    from catalyst_b200 import workloads
    ops = workloads.c3_qft_random(n, layers=3)
    d, err = _run(gpu, oracle_mod, n, ops)
    s = d.stats()
    assert err < TOL
    assert s.jit_passes == s.tile_passes > 0


@pytest.mark.skipif("B200SV_JIT" in __import__("os").environ, reason="the environment overrides the jit kwarg")
def test_jit_auto_mode_switches_over(gpu, oracle_mod):
    """'auto': the first sightings run interpreted, the structure is compiled in the background, later
    runs use the specialised kernels; the results never change beyond rounding."""
    from catalyst_b200 import workloads
    n = 16
    ops = workloads.c3_qft_random(n, layers=2)
    o = oracle_mod.OracleDevice(n)
    apply_ops(o, ops)
    ref = o.state()
    d = gpu.B200Device(n, jit="auto", jit_min_qubits=10, jit_threshold=2)
    seen = []
    for rep in range(4):
        d.set_basis_state([0] * n, list(range(n)))
        d.reset_stats()
        apply_ops(d, ops)
        assert np.max(np.abs(d.state() - ref)) < TOL
        seen.append(int(d.stats().jit_passes))
        d.jit_wait()
    assert seen[0] == 0 and seen[-1] > 0, seen


@pytest.mark.parametrize("absorber", [False, True])
@pytest.mark.parametrize("factor", [0, 1])
def test_jit_scale_factoring(gpu, oracle_mod, factor, absorber):
    """Uncontrolled one-qubit gates are divided by their u00 and the product of the scalars is folded
    into one op of the pass (simulator_jit.cpp factor_scales): angles next to pi (tiny and vanishing
    cos), Hadamard chains without and with a parity phase to absorb the product."""
    n = 13
    rng = np.random.default_rng(17)
    near = [np.pi, np.pi - 1e-3, np.pi + 3e-5, -np.pi + 1e-7, 0.0, 1e-9, 2 * np.pi, 3.0]
    ops = []
    for layer in range(6):
        for w in range(n):
            th = near[(layer * n + w) % len(near)] if layer % 2 == 0 else float(rng.uniform(-3, 3))
            ops.append((["RX", "RY", "Hadamard", "Rot"][(w + layer) % 4],
                        [th] if (w + layer) % 4 < 2 else ([] if (w + layer) % 4 == 2 else [th, 0.3, -1.1]), [w]))
        ops += [("CNOT", [], [w, (w + 1 + layer) % n]) for w in range(0, n, 2)]
        if absorber:
            ops.append(("RZ", [0.37 * (layer + 1)], [int(rng.integers(n))]))
            ops.append(("IsingZZ", [0.11], [0, n - 1]))
    d, err = _run(gpu, oracle_mod, n, ops, init=random_state(8, n), jit_factor=factor)
    s = d.stats()
    assert err < TOL
    assert s.jit_passes == s.tile_passes > 0


def test_jit_new_parameters_reuse_kernels(gpu, oracle_mod):
    n = 12
    rng = np.random.default_rng(5)
    d = gpu.B200Device(n, jit="always")
    compiled = []
    for rep in range(3):
        th = rng.uniform(-3, 3, size=3 * n)
        ops = [("RX", [th[w]], [w]) for w in range(n)] + [("CNOT", [], [w, (w + 1) % n]) for w in range(n)]
        ops += [("RZ", [th[n + w]], [w]) for w in range(n)] + [("IsingXY", [th[2 * n + w]], [w, (w + 5) % n]) for w in range(n)]
        o = oracle_mod.OracleDevice(n)
        apply_ops(o, ops)
        d.set_basis_state([0] * n, list(range(n)))
        apply_ops(d, ops)
        assert np.max(np.abs(d.state() - o.state())) < TOL
        compiled.append(int(d.stats().jit_compiled))
    assert compiled[1] == compiled[0] and compiled[2] == compiled[0], compiled


def test_jit_gradient_and_expval(gpu, oracle_mod):
    n = 12
    from catalyst_b200 import workloads
    ops, obs = workloads.c2_qaoa_maxcut(n=n, layers=2, seed=3)
    d = gpu.B200Device(n, jit="always")
    o = oracle_mod.OracleDevice(n)
    d.start_tape(); o.start_tape()
    apply_ops(d, ops); apply_ops(o, ops)
    ed, eo = d.expval(build_obs(d, obs)), o.expval(build_obs(o, obs))
    d.stop_tape(); o.stop_tape()
    assert abs(ed - eo) < 1e-12
    assert_close(d.gradient(), o.gradient(), "jit: adjoint gradient")
    assert d.stats().jit_passes > 0


def test_jit_sharded_virtual_ranks(gpu, oracle_mod):
    n, world = 12, 4
    ops = random_circuit(21, n, 60, names=NO_DENSE)
    o = oracle_mod.OracleDevice(n)
    apply_ops(o, ops)
    ref = o.state()

    def body(d, r):
        d.allocate_qubits(n)
        apply_ops(d, ops)
        return d.state(), int(d.stats().jit_passes)

    res = gpu.run_spmd(world, body, jit="always")
    for st, jp in res:
        assert np.max(np.abs(st - ref)) < TOL
        assert jp > 0


# ---- tile-uniform diagonal factors (UAcc in simulator_jit.cpp) -----------------------------------------------
DIAG_MIX = ["ControlledPhaseShift", "CZ", "RZ", "IsingZZ", "MultiRZ", "PhaseShift", "T", "S", "CRZ", "PCPhase",
            "GlobalPhase", "Hadamard", "RX", "CNOT"]


@pytest.mark.parametrize("uniform", [0, 1])
@pytest.mark.parametrize("n,tile_bits,seed", [(14, 8, 1), (17, 9, 2), (20, 11, 3)])
def test_jit_uniform_diagonal_factors(gpu, oracle_mod, n, tile_bits, seed, uniform):
    """Long fans of diagonal ops whose predicates / parities sit on bits OUTSIDE the tile (incl. negative control
    values, parity over several outside bits, controlled parities): evaluated once per tile by a lane group
    (jit_uniform=1) they must equal the per-thread evaluation (jit_uniform=0) and the oracle."""
    rng = np.random.default_rng(seed)
    ops = []
    for j in range(0, n, 3):                                   # QFT-like fans: H(j) then phases controlled on every other wire
        ops.append(("Hadamard", [], [j]))
        for k in range(n):
            if k != j:
                ops.append(("ControlledPhaseShift", [float(rng.uniform(0, 6))], [k, j]))
        for k in range(0, n, 2):
            if k != j:
                ops.append(("PhaseShift", [float(rng.uniform(0, 6))], [j],
                            {"ctrl_wires": [k], "ctrl_values": [bool(rng.integers(2))]}))
        ws = [int(w) for w in rng.permutation(n)[:4]]
        ops.append(("MultiRZ", [float(rng.uniform(0, 6))], ws))
        ops.append(("IsingZZ", [float(rng.uniform(0, 6))], ws[:2]))
        ops.append(("RZ", [float(rng.uniform(0, 6))], [ws[2]], {"ctrl_wires": [ws[3]], "ctrl_values": [True]}))
    ops += random_circuit(seed, n, 150, names=DIAG_MIX)
    d, err = _run(gpu, oracle_mod, n, ops, init=random_state(seed, n), jit_uniform=uniform, tile_bits=tile_bits,
                  jit_regbits=3 if tile_bits == 8 else 4)
    s = d.stats()
    assert s.jit_passes == s.tile_passes > 0
    assert err < TOL, err


def test_jit_disk_cache_across_processes(tmp_path):
    """A second PROCESS that meets the same pass structures loads the cubins from $B200SV_JIT_CACHE at the first
    sighting: nothing is compiled and every pass of its FIRST run is specialised.  (Each run ends with State(),
    which turns the wire relabelling of the QFT's SWAPs back into real swaps fused into the last passes: those
    structures are part of the run.)  The process also exits cleanly right after submitting work."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    prog = (
        "import sys, json; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from catalyst_b200 import capi, workloads\nfrom dsl import apply_ops\n"
        "d = capi.B200Device(16, jit_min_qubits=10, jit_threshold=1)\n"
        "runs = []\n"
        "for rep in range(2):\n"
        "    d.set_basis_state([0]*16, list(range(16))); d.reset_stats()\n"
        "    apply_ops(d, workloads.c3_qft_random(16, layers=3)); st = d.state(); s = d.stats()\n"
        "    runs.append([int(s.jit_passes), int(s.tile_passes)]); d.jit_wait()\n"
        "s = d.stats(); d.close()\n"
        "print(json.dumps({'compiled': s.jit_compiled, 'disk': s.jit_disk_hits, 'runs': runs,\n"
        "                  'norm': float(abs((st.conj()*st).sum())), 'a0': [st[5].real, st[5].imag]}))\n"
    ) % (root, os.path.join(root, "tests"))
    env = dict(os.environ, B200SV_JIT_CACHE=str(tmp_path / "cache"), B200SV_JIT="auto", B200SV_DEBUG_SEGV="1")
    out = []
    for _ in range(2):
        r = subprocess.run([sys.executable, "-c", prog], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (r.stdout[-600:], r.stderr[-3000:])
        out.append(json.loads(r.stdout.strip().splitlines()[-1]))
    first, second = out
    assert first["compiled"] > 0 and first["disk"] == 0
    assert first["runs"][0][0] == 0 and first["runs"][1][0] == first["runs"][1][1] > 0   # interpreted, then specialised
    assert second["compiled"] == 0 and second["disk"] == first["compiled"]
    assert second["runs"][0][0] == second["runs"][0][1] > 0, "the second process must run specialised from its first pass"
    assert abs(second["norm"] - 1.0) < 1e-12 and second["a0"] == first["a0"]


def test_process_exit_with_a_compilation_in_flight():
    """A short-lived process whose last call submits a NEW pass structure to the background compiler and then
    exits at once must not crash: the last device going away cancels queued compilations and waits for the
    running one (seen on the GPU box as SIGSEGV inside nvrtcCompileProgram during exit-time destructors)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    prog = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from catalyst_b200 import capi, workloads\nfrom dsl import apply_ops\n"
        "d = capi.B200Device(16, jit_min_qubits=10, jit_threshold=1)\n"
        "apply_ops(d, workloads.c3_qft_random(16, layers=3)); d.synchronize()\n"
        "print('submitted', d.stats().jit_pending, flush=True)\n"
    ) % (root, os.path.join(root, "tests"))
    env = dict(os.environ, B200SV_JIT_CACHE="off", B200SV_JIT="auto", B200SV_DEBUG_SEGV="1")
    for _ in range(3):
        r = subprocess.run([sys.executable, "-c", prog], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, (r.stdout[-300:], r.stderr[-3000:])
        assert "submitted" in r.stdout
