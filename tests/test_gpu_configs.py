"""Parity of the CUDA device on BASELINE.json's five configs at (or near) their sizes.

Two kinds of check per config, both through the C ABI (catalyst_b200.capi -> libb200sv.so):

* LIVE: the exact `workloads.cN` generator runs on the CUDA device and on the CPU oracle in the same
  process, at the largest register the oracle finishes in well under a minute on the GPU box's host.
* FIXTURE: the same program at a larger register against tests/golden/configs/*.npz, which
  scripts/gen_config_goldens.py produced with the same oracle in the build container (minutes of CPU).

Tolerance: 1e-12 RELATIVE (max |device - oracle| / max |oracle|) for amplitudes, probabilities and gradients,
and |device - oracle| / max(|oracle|, ||O||) for expectation values of Pauli words (||O|| = 1: the <Z0> of a
deep random circuit is ~1e-4, its rounding noise ~1e-16), as BASELINE.json's north_star states; sampled basis states and
mid-circuit measurement outcomes bit-exact under the seeded stream.  Every test prints the error it
achieved (run with -s, or see profiles/r2_*_pytest_gpu.log).
"""
import os

import numpy as np
import pytest

import config_cases as cc
from catalyst_b200 import capi

pytestmark = pytest.mark.gpu

TOL = 1e-12
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "configs")


def fixture(name):
    hits = sorted(f for f in os.listdir(GOLD) if f.startswith(name + "_n")) if os.path.isdir(GOLD) else []
    if not hits:
        pytest.skip(f"no golden fixture for {name} (scripts/gen_config_goldens.py)")
    z = np.load(os.path.join(GOLD, hits[-1]))
    return int(z["n"]), z


def report(tag, **errs):
    print(f"[parity] {tag}: " + ", ".join(f"{k} rel err {v:.3e}" for k, v in errs.items()))
    for k, v in errs.items():
        assert v <= TOL, f"{tag}: {k} relative error {v:.3e} exceeds {TOL}"


# ---- C1: 20-qubit StronglyEntanglingLayers (4 layers), expval(PauliZ(0)) -----------------------------
@pytest.mark.parametrize("jit", ["off", "always"])
def test_c1_full_size_live(oracle_mod, jit):
    n = 20
    ref = cc.run_c1(oracle_mod.OracleDevice(n, seed=cc.SEED), n)
    d = capi.B200Device(n, seed=cc.SEED, jit=jit)
    got = cc.run_c1(d, n)
    d.close()
    report(f"C1 n={n} jit={jit}", expval=cc.expval_err(got["expval"], ref["expval"]),
           state=cc.rel_err(got["state"], ref["state"]))


# ---- C2: QAOA MaxCut, 6 layers, adjoint gradient -------------------------------------------------------
def test_c2_live(oracle_mod):
    n = 20
    ref = cc.run_c2(oracle_mod.OracleDevice(n, seed=cc.SEED), n)
    d = capi.B200Device(n, seed=cc.SEED, jit="always")
    got = cc.run_c2(d, n)
    d.close()
    assert got["grad"].shape == ref["grad"].shape == (1, 6 * (3 * n // 2 + n))
    report(f"C2 n={n}", expval=cc.rel_err(got["expval"], ref["expval"]), grad=cc.rel_err(got["grad"], ref["grad"]),
           state=cc.rel_err(got["state"], ref["state"]))


def test_c2_fixture():
    n, z = fixture("c2")
    d = capi.B200Device(n, seed=cc.SEED)
    got = cc.run_c2(d, n)
    d.close()
    report(f"C2 n={n} (fixture)", expval=cc.rel_err(got["expval"], z["expval"]),
           grad=cc.rel_err(got["grad"], z["grad"]), state=cc.rel_err(got["state_probe"], z["state_probe"]))


# ---- C3: QFT + random layers, probs() and 1e5 shots -----------------------------------------------------
def test_c3_live(oracle_mod):
    n = 24
    ref = cc.run_c3(oracle_mod.OracleDevice(n, seed=cc.SEED), n)
    # jit='auto': the first two runs are interpreted, the third runs on the specialised kernels
    d = capi.B200Device(n, seed=cc.SEED, jit="auto")
    got = cc.run_c3(d, n, repeats=3)
    s = d.stats()
    d.close()
    assert s.jit_passes > 0, "the third run was expected on specialised passes"
    report(f"C3 n={n} probs", probs=cc.rel_err(got["probs"], ref["probs"]), mass=abs(got["probs_sum"] - 1.0))
    # repeats=3 took no extra draws from the device generator: the sampled stream must be the oracle's
    assert np.array_equal(got["sample_idx"], ref["sample_idx"]), "C3: 1e5 seeded samples differ from the oracle"
    print(f"[parity] C3 n={n}: {cc.SHOTS} samples bit-exact")


def test_c3_fixture():
    n, z = fixture("c3")
    d = capi.B200Device(n, seed=cc.SEED)
    got = cc.run_c3(d, n)
    d.close()
    report(f"C3 n={n} (fixture)", probs=cc.rel_err(got["probs_probe"], z["probs_probe"]),
           mass=abs(got["probs_sum"] - float(z["probs_sum"])))
    assert np.array_equal(got["sample_head"], z["sample_head"])
    assert got["sample_sha256"] == str(z["sample_sha256"]), "C3: SHA-256 of 1e5 sampled basis states differs"
    print(f"[parity] C3 n={n}: {cc.SHOTS} samples bit-exact (SHA-256)")


# ---- C4: VQE, 1000-term Hamiltonian expval + adjoint Jacobian + mid-circuit measurement ------------------
def check_c4(tag, got, ref):
    assert np.array_equal(got["mcm_outcomes"], ref["mcm_outcomes"]), (got["mcm_outcomes"], ref["mcm_outcomes"])
    report(tag, expval=cc.rel_err(got["expval"], ref["expval"]), grad=cc.rel_err(got["grad"], ref["grad"]),
           mcm_expvals=cc.rel_err(got["mcm_expvals"], ref["mcm_expvals"]),
           mcm_probs=cc.rel_err(got["mcm_probs"], ref["mcm_probs"]))


def test_c4_live(oracle_mod):
    n = 18
    ref = cc.run_c4(oracle_mod.OracleDevice(n, seed=cc.SEED), n)
    d = capi.B200Device(n, seed=cc.SEED, jit="always")
    got = cc.run_c4(d, n)
    d.close()
    assert got["grad"].shape == (1, 4 * 2 * n)
    check_c4(f"C4 n={n} 1000 terms", got, ref)


def test_c4_fixture():
    n, z = fixture("c4")
    d = capi.B200Device(n, seed=cc.SEED)
    got = cc.run_c4(d, n)
    d.close()
    check_c4(f"C4 n={n} 1000 terms (fixture)", got, z)


# ---- C5: random circuit (depth 40) on a sharded register --------------------------------------------------
_c5_ref = {}


def c5_reference(oracle_mod, n, depth):
    if (n, depth) not in _c5_ref:
        _c5_ref[(n, depth)] = cc.run_c5(oracle_mod.OracleDevice(n, seed=cc.SEED), n, depth)
    return _c5_ref[(n, depth)]


@pytest.mark.parametrize("world,p2p", [(2, -1), (4, 0), (8, -1), (8, 0)])
def test_c5_virtual_shards_live(oracle_mod, world, p2p):
    n, depth = 22, 40
    ref = c5_reference(oracle_mod, n, depth)
    res = capi.run_spmd(world, lambda d, r: (d.allocate_qubits(n), cc.run_c5(d, n, depth), d.stats().exchanges)[1:],
                        seed=cc.SEED, p2p_swap=p2p, jit="always" if world == 8 else "auto", jit_min_qubits=10)
    assert res[0][1] > 0, "the circuit was expected to need global<->local qubit swaps"
    for r, (got, _) in enumerate(res):
        report(f"C5 n={n} depth={depth} world={world} p2p={p2p} rank {r}",
               expval=cc.expval_err(got["expval"], ref["expval"]), expval_zx=cc.expval_err(got["expval_zz"], ref["expval_zz"]),
               partial_probs=cc.rel_err(got["partial_probs"], ref["partial_probs"]),
               state=cc.rel_err(got["state"], ref["state"]))


@pytest.mark.parametrize("world", [8])
def test_c5_virtual_shards_fixture(world):
    n, z = fixture("c5")
    res = capi.run_spmd(world, lambda d, r: (d.allocate_qubits(n), cc.run_c5(d, n, 40, state=False))[1], seed=cc.SEED)
    for r, got in enumerate(res):
        report(f"C5 n={n} depth=40 world={world} rank {r} (fixture)", expval=cc.expval_err(got["expval"], z["expval"]),
               expval_zx=cc.expval_err(got["expval_zz"], z["expval_zz"]),
               partial_probs=cc.rel_err(got["partial_probs"], z["partial_probs"]))


def test_c5_single_gpu_fixture():
    """The same 26-qubit program unsharded: amplitudes at the probe indices against the fixture."""
    n, z = fixture("c5")
    d = capi.B200Device(n, seed=cc.SEED)
    got = cc.run_c5(d, n, 40)
    d.close()
    report(f"C5 n={n} depth=40 single GPU (fixture)", expval=cc.expval_err(got["expval"], z["expval"]),
           partial_probs=cc.rel_err(got["partial_probs"], z["partial_probs"]),
           state=cc.rel_err(got["state_probe"], z["state_probe"]))


# ---- the planner's refinement step (automatic from 2^27 local amplitudes) forced on at test sizes ---------------------
def test_planner_refinement_forced_c3(oracle_mod):
    n = 22
    ref = cc.run_c3(oracle_mod.OracleDevice(n, seed=cc.SEED), n)
    d = capi.B200Device(n, seed=cc.SEED, jit="always", plan_refine=6)
    got = cc.run_c3(d, n)
    d.close()
    report(f"C3 n={n} plan_refine=6", probs=cc.rel_err(got["probs"], ref["probs"]))
    assert np.array_equal(got["sample_idx"], ref["sample_idx"])


@pytest.mark.parametrize("world", [1, 4])
def test_planner_refinement_forced_c5(oracle_mod, world):
    n, depth = 22, 40
    ref = c5_reference(oracle_mod, n, depth)
    res = capi.run_spmd(world, lambda d, r: (d.allocate_qubits(n), cc.run_c5(d, n, depth))[1], seed=cc.SEED,
                        plan_refine=6, jit_min_qubits=10)
    for r, got in enumerate(res):
        report(f"C5 n={n} world={world} plan_refine=6 rank {r}", expval=cc.expval_err(got["expval"], ref["expval"]),
               partial_probs=cc.rel_err(got["partial_probs"], ref["partial_probs"]), state=cc.rel_err(got["state"], ref["state"]))
