"""The NCCL transport of the sharded device: one PROCESS per GPU (the torchrun model), ncclUniqueId
handed over by the parent.  Needs >= 2 GPUs; on a single-GPU box these tests skip and the same
logic is covered by tests/test_gpu_sharded.py (threads + 'local:' group)."""
import multiprocessing as mp
import os
import sys

import numpy as np
import pytest

from dsl import apply_ops, build_obs, random_circuit, random_pauli_obs

pytestmark = pytest.mark.gpu


def _ngpus():
    try:
        import subprocess
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def _worker(rank, world, uid, n, ops, obs_list, q, kwargs=None):
    try:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        sys.path[:0] = [root, os.path.join(root, "tests")]
        from catalyst_b200 import capi
        d = capi.B200Device(world_size=world, rank=rank, comm="nccl:" + uid, device=rank, seed=37,
                            **(kwargs or {}))
        d.allocate_qubits(n)
        apply_ops(d, ops)
        e = [d.expval(build_obs(d, ob)) for ob in obs_list]
        p = d.probs([0, n - 1, 3])
        m = d.measure(0)
        smp = d.sample(200, [1, 2, n - 1])
        st = d.state()
        s = d.stats()
        q.put((rank, e, p, m, smp, st, (s.exchanges, s.swap_push_exchanges, s.p2p_exchanges), None))
        d.close()
    except BaseException as ex:  # noqa: BLE001
        q.put((rank, None, None, None, None, None, (0, 0, 0), repr(ex)))


@pytest.mark.parametrize("world,n,kwargs", [
    # automatic policy: push mode in a world of 2 ranks, the staged NCCL path at 4 and 8
    (2, 14, None), (4, 14, None), (8, 14, None),
    # push mode forced (one kernel writes into every peer's staging, mapped through CUDA IPC) at 4 and 8 ranks
    (4, 14, {"swap_push": 1}), (8, 14, {"swap_push": 1}),
    # 1 MiB staging at 18 qubits: every swap runs as several chunks -- push mode (both staging halves, the trailing
    # unpack-only round) and the staged NCCL path (pack k+1 / ncclSend+ncclRecv k / unpack k-1)
    (2, 18, {"stage_mib": 1, "swap_push": 1}), (4, 18, {"stage_mib": 1, "swap_push": 1}), (8, 18, {"stage_mib": 1, "swap_push": 1}),
    (2, 18, {"stage_mib": 1, "p2p_swap": 0, "swap_push": 0}), (4, 18, {"stage_mib": 1, "p2p_swap": 0, "swap_push": 0}),
    (8, 16, {"p2p_swap": 0, "swap_push": 0}),
    # the in-place pairwise kernel over the peers' mapped shards
    (2, 14, {"p2p_swap": 1, "swap_push": 0}), (8, 14, {"p2p_swap": 1, "swap_push": 0})])
def test_nccl_ranks_match_oracle(oracle_mod, world, n, kwargs):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    from catalyst_b200 import capi
    ops = random_circuit(500 + world, n, 150)
    rng = np.random.default_rng(3)
    obs_list = [random_pauli_obs(rng, n) for _ in range(4)]
    o = oracle_mod.OracleDevice(n, seed=37)
    apply_ops(o, ops)
    want_e = [o.expval(build_obs(o, ob)) for ob in obs_list]
    want_p = o.probs([0, n - 1, 3])
    want_m = o.measure(0)
    want_s = o.sample(200, [1, 2, n - 1])
    want_st = o.state()
    uid = capi.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, uid, n, ops, obs_list, q, kwargs)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, e, p, m, smp, st, nx, err in res:
        assert err is None, f"rank {rank}: {err}"
        assert np.max(np.abs(np.array(e) - want_e)) < 1e-12
        assert np.max(np.abs(p - want_p)) < 1e-12
        assert m == want_m
        assert np.array_equal(smp, want_s)
        assert np.max(np.abs(st - want_st)) < 1e-12
        assert nx[0] > 0
        kw = kwargs or {}
        if kw.get("swap_push", -1) == 1 or (not kw and world == 2):
            assert nx[1] == nx[0], f"expected every swap in push mode, got {nx}"
        if kw.get("swap_push", -1) == 0:
            assert nx[1] == 0
        print(f"[nccl parity] world={world} n={n} {kw}: swaps (all, push, in-place) = {nx}, "
              f"max state err {np.max(np.abs(st - want_st)):.2e}")


def _worker_sampler(rank, world, uid, n, ops, q):
    try:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        sys.path[:0] = [root, os.path.join(root, "tests")]
        from catalyst_b200 import capi
        # dist_sampler=1: the distributed sampler is REQUIRED (the slices of |psi|^2 are read through CUDA IPC mappings)
        d = capi.B200Device(world_size=world, rank=rank, comm="nccl:" + uid, device=rank, seed=37, dist_sampler=1,
                            gather="slice")
        d.allocate_qubits(n)
        apply_ops(d, ops)
        smp = d.sample(3000)
        cnt = d.counts(500)
        smp2 = d.sample(100, [2, n - 1])
        st = np.full(1 << n, np.nan + 0j, dtype=np.complex128)
        d.state(out=st)
        p = np.full(1 << n, np.nan)
        d.probs(out=p)
        q.put((rank, smp, cnt, smp2, st, p, None))
        d.close()
    except BaseException as ex:  # noqa: BLE001
        q.put((rank, None, None, None, None, None, repr(ex)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_distributed_sampler_and_slices(oracle_mod, world):
    """Sample / Counts of all wires through the distributed sampler (peer reads of the ranks' |psi|^2 slices) and
    State / Probs in gather='slice' mode, one process per GPU over NCCL: bit-exact samples, 1e-12 slices."""
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    from catalyst_b200 import capi
    n = 16
    nl = n - (world.bit_length() - 1)
    ops = random_circuit(900 + world, n, 150)
    o = oracle_mod.OracleDevice(n, seed=37)
    apply_ops(o, ops)
    want_s, want_c, want_s2 = o.sample(3000), o.counts(500), o.sample(100, [2, n - 1])
    want_st, want_p = o.state(), o.probs()
    uid = capi.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_sampler, args=(r, world, uid, n, ops, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, smp, cnt, smp2, st, p, err in res:
        assert err is None, f"rank {rank}: {err}"
        assert np.array_equal(smp, want_s)
        assert np.array_equal(cnt[0], want_c[0]) and np.array_equal(cnt[1], want_c[1])
        assert np.array_equal(smp2, want_s2)
        lo, hi = rank << nl, (rank + 1) << nl
        assert np.max(np.abs(st[lo:hi] - want_st[lo:hi])) < 1e-12
        assert np.max(np.abs(p[lo:hi] - want_p[lo:hi])) < 1e-12
        assert np.all(np.isnan(p[:lo])) and np.all(np.isnan(p[hi:]))
    print(f"[nccl parity] world={world} n={n}: distributed sampler bit-exact (3000 + 500 + 100 shots), slices 1e-12")
