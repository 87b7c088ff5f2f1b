"""Synthetic workloads of BASELINE.json's five configs, as op lists in the tests'
mini-DSL: (name, params, wires[, kwargs]).  Pure host-side circuit descriptions — no numerics.
Shared by bench.py, the GPU parity tests and the CPU-baseline leg so that every arm runs exactly
the same gate sequence."""
from __future__ import annotations

import numpy as np


def _angles(rng, k):
    return [float(x) for x in rng.uniform(0.0, 2.0 * np.pi, size=k)]


def c1_strongly_entangling(n=20, layers=4, seed=0):
    """Config 1: StronglyEntanglingLayers — per layer n x Rot then n x CNOT(i, (i+r) % n),
    r = (layer % (n-1)) + 1; observable PauliZ(0)."""
    rng = np.random.default_rng(seed)
    ops = []
    for layer in range(layers):
        for w in range(n):
            ops.append(("Rot", _angles(rng, 3), [w]))
        r = (layer % (n - 1)) + 1 if n > 1 else 0
        if n > 1:
            for w in range(n):
                ops.append(("CNOT", [], [w, (w + r) % n]))
    return ops, ("N", "PauliZ", 0)


def _regular_graph(n, degree, rng):
    """Random d-regular graph by repeated pairing (retry on loops / multi-edges)."""
    assert (n * degree) % 2 == 0
    while True:
        stubs = np.repeat(np.arange(n), degree)
        rng.shuffle(stubs)
        edges = set()
        ok = True
        for a, b in zip(stubs[0::2], stubs[1::2]):
            a, b = int(min(a, b)), int(max(a, b))
            if a == b or (a, b) in edges:
                ok = False
                break
            edges.add((a, b))
        if ok:
            return sorted(edges)


def c2_qaoa_maxcut(n=28, layers=6, seed=0, two_qubit="IsingZZ"):
    """Config 2: QAOA MaxCut on a random 3-regular graph: n Hadamards, then per layer one
    IsingZZ(2*gamma) per edge and RX(2*beta) per wire; cost Hamiltonian sum 0.5 * Z_i Z_j."""
    rng = np.random.default_rng(seed)
    edges = _regular_graph(n, 3, rng)
    ops = [("Hadamard", [], [w]) for w in range(n)]
    for _ in range(layers):
        gamma, beta = _angles(rng, 2)
        for (a, b) in edges:
            ops.append((two_qubit, [2 * gamma], [a, b]))
        for w in range(n):
            ops.append(("RX", [2 * beta], [w]))
    ham = ("Ham", [0.5] * len(edges), [("T", [("N", "PauliZ", a), ("N", "PauliZ", b)]) for a, b in edges])
    return ops, ham


def qft(n):
    ops = []
    for i in range(n):
        ops.append(("Hadamard", [], [i]))
        for j in range(i + 1, n):
            ops.append(("ControlledPhaseShift", [float(np.pi / 2 ** (j - i))], [j, i]))
    for i in range(n // 2):
        ops.append(("SWAP", [], [i, n - 1 - i]))
    return ops


def random_layers(n, layers, seed, one_q=("RX", "RY", "RZ", "Hadamard", "T"), two_q=("CNOT",)):
    rng = np.random.default_rng(seed)
    ops = []
    for _ in range(layers):
        for w in range(n):
            g = one_q[int(rng.integers(len(one_q)))]
            ops.append((g, _angles(rng, 1) if g in ("RX", "RY", "RZ") else [], [w]))
        perm = [int(x) for x in rng.permutation(n)]
        for a, b in zip(perm[0::2], perm[1::2]):
            ops.append((two_q[int(rng.integers(len(two_q)))], [], [a, b]))
    return ops


def c3_qft_random(n=30, layers=10, seed=0):
    """Config 3: QFT followed by `layers` layers of n random one-qubit gates and a random perfect
    matching of CNOTs; measured with probs() and 1e5 shots."""
    return qft(n) + random_layers(n, layers, seed)


def c4_vqe(n=26, layers=4, terms=1000, seed=0, p_identity=0.85):
    """Config 4: hardware-efficient ansatz (RY, RZ per wire + CNOT chain per layer) and a random
    `terms`-term Pauli Hamiltonian."""
    rng = np.random.default_rng(seed)
    ops = []
    for _ in range(layers):
        for w in range(n):
            ops.append(("RY", _angles(rng, 1), [w]))
        for w in range(n):
            ops.append(("RZ", _angles(rng, 1), [w]))
        for w in range(n - 1):
            ops.append(("CNOT", [], [w, w + 1]))
    names = {1: "PauliX", 2: "PauliY", 3: "PauliZ"}
    obs_terms = []
    for _ in range(terms):
        while True:
            word = rng.choice(4, size=n, p=[p_identity] + [(1 - p_identity) / 3] * 3)
            if word.any():
                break
        fs = [("N", names[int(c)], int(w)) for w, c in enumerate(word) if c]
        obs_terms.append(fs[0] if len(fs) == 1 else ("T", fs))
    coeffs = [float(x) for x in rng.normal(size=terms)]
    return ops, ("Ham", coeffs, obs_terms)


def c5_random_depth(n=34, depth=40, seed=0):
    """Config 5: `depth` layers of n random {RX, RY, RZ} and n/2 CZ/CNOT on a random matching over
    all wires (so some two-qubit gates touch the sharded high-order qubits)."""
    return random_layers(n, depth, seed, one_q=("RX", "RY", "RZ"), two_q=("CZ", "CNOT"))


def gate_count(ops):
    return sum(1 for op in ops if op[0] not in ("Identity",))
