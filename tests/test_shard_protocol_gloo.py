"""World-size-2 CPU tests (gloo) of the host-side rules the sharded device follows (DESIGN.md section 5;
C++ implementation: catalyst_b200/csrc/simulator.cpp `swap_bits` / `execute`):

* rank-bit layout: rank r owns the amplitudes whose top index bit is r, `pos[q]` = physical position of
  logical bit q;
* control wires and diagonal targets on the rank bit need no data movement (predicate / phase chosen by r);
* a non-diagonal target on the rank bit triggers the one exchange step: rank a trades its sub-block whose
  chosen local bit spells b with rank b's sub-block that spells a, then the two logical qubits trade places;
* reductions: per-rank partials are all-gathered and summed in rank order, so every rank holds the same bits.

Each rank keeps its shard as a numpy array and moves data with torch.distributed send/recv over gloo; the
2x2 matrices come from the CPU oracle and the assembled state is compared with the oracle's on the same
circuit.  This is a model of the protocol, not the CUDA code: the CUDA code runs the same circuits in
tests/test_gpu_sharded.py (threads) and tests/test_gpu_nccl.py (processes + NCCL).  The same two processes
also exercise bench.py's control plane (`Plumbing`: barrier / max / sum / broadcast of the NCCL id)."""
import multiprocessing as mp
import os
import socket
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DIAGONAL = {"RZ", "T", "PauliZ", "S", "PhaseShift"}


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _apply_local(shard, u, p, ctrl_bits):
    """2x2 matrix u on local bit p of the shard, on the amplitudes whose `ctrl_bits` are all 1."""
    i0 = np.arange(shard.size)
    keep = (i0 >> p) & 1 == 0
    for c in ctrl_bits:
        keep &= (i0 >> c) & 1 == 1
    i0 = i0[keep]
    i1 = i0 | (1 << p)
    a0, a1 = shard[i0].copy(), shard[i1].copy()
    shard[i0] = u[0, 0] * a0 + u[0, 1] * a1
    shard[i1] = u[1, 0] * a0 + u[1, 1] * a1


def _worker(rank, world, port, n, ops, mats, q):
    try:
        sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch
        import torch.distributed as dist
        import bench
        plumb = bench.Plumbing(world)          # init_process_group("gloo") from the env, as under torchrun
        assert plumb.max(float(rank)) == world - 1 and plumb.sum(1.0) == world
        uid = plumb.bcast_obj("ab" * 64 if rank == 0 else None)
        plumb.barrier()

        nl = n - 1                              # local bits; physical position nl is the rank bit
        pos = list(range(n))                    # pos[q]: physical position of logical bit q (= n-1-wire)
        shard = np.zeros(1 << nl, dtype=np.complex128)
        if rank == 0:
            shard[0] = 1.0
        swaps = 0

        def swap_rank_bit_with(p):
            nonlocal swaps
            peer = 1 - rank
            idx = np.arange(shard.size)
            sel = idx[(idx >> p) & 1 == peer]   # my sub-block that spells the peer's pattern
            out = torch.from_numpy(np.ascontiguousarray(shard[sel]).view(np.float64))
            inc = torch.empty_like(out)
            reqs = [dist.isend(out, peer), dist.irecv(inc, peer)]
            for r in reqs:
                r.wait()
            shard[sel] = inc.numpy().view(np.complex128)
            a, b = pos.index(p), pos.index(nl)
            pos[a], pos[b] = nl, p
            swaps += 1

        for name, params, wires in ops:
            ctrl = []
            if name in ("CNOT", "CZ"):
                ctrl, tgt, base = [n - 1 - wires[0]], n - 1 - wires[1], ("PauliX" if name == "CNOT" else "PauliZ")
                u = mats[(base, ())]
            else:
                tgt, base = n - 1 - wires[0], name
                u = mats[(name, tuple(params))]
            if pos[tgt] == nl and base not in DIAGONAL:
                busy = {pos[c] for c in ctrl}
                swap_rank_bit_with(min(p for p in range(nl) if p not in busy))
            cb = [pos[c] for c in ctrl]
            if nl in cb:                        # control on the rank bit: a predicate on r
                if rank == 0:
                    continue
                cb.remove(nl)
            if pos[tgt] == nl:                  # diagonal target on the rank bit: a phase chosen by r
                i = np.arange(shard.size)
                keep = np.ones(shard.size, dtype=bool)
                for c in cb:
                    keep &= (i >> c) & 1 == 1
                shard[keep] *= u[rank, rank]
            else:
                _apply_local(shard, u, pos[tgt], cb)

        # <Z> on every wire: per-rank partials, all-gathered, summed in rank order on every rank
        ez = []
        for w in range(n):
            p = pos[n - 1 - w]
            if p == nl:
                part = (1.0 - 2.0 * rank) * float(np.sum(np.abs(shard) ** 2))
            else:
                sign = 1.0 - 2.0 * ((np.arange(shard.size) >> p) & 1)
                part = float(np.sum(sign * np.abs(shard) ** 2))
            box = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
            dist.all_gather(box, torch.tensor([part], dtype=torch.float64))
            total = 0.0
            for t in box:
                total += float(t.item())
            ez.append(total)
        # State: every rank's slice, layout undone on the host
        parts = [torch.zeros(2 << nl, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(shard.view(np.float64).copy()))
        phys = np.concatenate([t.numpy().view(np.complex128) for t in parts])
        L = np.arange(1 << n)
        P = np.zeros_like(L)
        for qb in range(n):
            P |= ((L >> qb) & 1) << pos[qb]
        state = phys[P]
        plumb.close()
        q.put((rank, uid, ez, state, swaps, None))
    except BaseException as ex:  # noqa: BLE001
        q.put((rank, None, None, None, 0, repr(ex)))


def test_world2_shard_protocol_matches_oracle(oracle_mod):
    n, world = 7, 2
    rng = np.random.default_rng(11)
    ops = []
    for _ in range(6):
        for w in range(n):
            g = ["Hadamard", "RX", "RY", "RZ", "T"][rng.integers(5)]
            ops.append((g, [float(rng.uniform(0, 2 * np.pi))] if g in ("RX", "RY", "RZ") else [], [w]))
        perm = rng.permutation(n)
        for k in range(0, n - 1, 2):
            ops.append((["CNOT", "CZ"][rng.integers(2)], [], [int(perm[k]), int(perm[k + 1])]))
    # gate matrices from the oracle: column j = the gate applied to |j>
    mats = {}
    for name, params, _ in ops + [("PauliX", [], [0]), ("PauliZ", [], [0])]:
        if name in ("CNOT", "CZ") or (name, tuple(params)) in mats:
            continue
        cols = []
        for j in range(2):
            o = oracle_mod.OracleDevice(1)
            if j:
                o.named_op("PauliX", [], [0])
            o.named_op(name, params, [0])
            cols.append(o.state())
        mats[(name, tuple(params))] = np.stack(cols, axis=1)
    o = oracle_mod.OracleDevice(n)
    for name, params, wires in ops:
        o.named_op(name, params, wires)
    want = o.state()
    want_ez = [o.expval(o.obs_named("PauliZ", w)) for w in range(n)]

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, ops, mats, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
    for rank, uid, ez, state, swaps, err in res:
        assert err is None, f"rank {rank}: {err}"
        assert uid == "ab" * 64                      # the NCCL id reaches every rank unchanged
        assert swaps >= 2                            # the circuit does need the exchange step
        assert np.max(np.abs(state - want)) < 1e-12
        assert np.max(np.abs(np.array(ez) - want_ez)) < 1e-12
    assert res[0][2] == res[1][2]                    # rank-order sums: bit-identical scalars on every rank
    assert res[0][4] == res[1][4]


def test_reference_arm_runs_on_rank0_only(capsys):
    """bench.py --impl reference under torchrun: ranks other than 0 leave without work or output."""
    sys.path[:0] = [ROOT]
    import bench
    assert bench.run_reference(None, 1, 2) is None
    assert capsys.readouterr().out == ""


# ---- the distributed sampler's protocol (C++: Simulator::sample_all_distributed, kernels
# k_samp_chunk_sums / k_samp_seg_* / k_samp_draw<SampSharded>) as a numpy model on two gloo ranks -------------------
SAMP_CH = 1024


def _mt19937_u32(seed, count):
    """Raw 32-bit outputs of std::mt19937(seed) (numpy's legacy seeding is init_genrand)."""
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed))
    return bg.random_raw(count).astype(np.uint64)


def _uniform_doubles(seed, count):
    """std::uniform_real_distribution<double>(0, 1) over std::mt19937(seed) as libstdc++ computes it
    (generate_canonical<double, 53>: two engine outputs per double, low word first)."""
    x = _mt19937_u32(seed, 2 * count).astype(np.float64)
    u = (x[0::2] + x[1::2] * 4294967296.0) / 18446744073709551616.0
    return np.where(u >= 1.0, np.nextafter(1.0, 0.0), u)


def _chunk_sums(p_slice, n_total):
    """Deficit / excess sums of the slice's 1024-entry chunks, each summed sequentially in DESCENDING index order."""
    n = len(p_slice)
    nch = n // SAMP_CH
    D, E = np.zeros(nch), np.zeros(nch)
    for q in range(nch):
        d = e = 0.0
        for i in range(n - 1 - q * SAMP_CH, n - 1 - (q + 1) * SAMP_CH, -1):
            f = n_total * p_slice[i]
            if f < 1.0:
                d += 1.0 - f
            else:
                e += f - 1.0
        D[q], E[q] = d, e
    return D, E


def _exclusive(sums):
    cum = np.zeros(len(sums) + 1)
    run = 0.0
    for q, v in enumerate(sums):
        cum[q] = run
        run += v
    cum[len(sums)] = run
    return cum


def _find_chunk(cum, nch, x, strict):
    for q in range(nch):
        if (cum[q + 1] > x) if strict else (cum[q + 1] >= x):
            return q
    return nch


def _draw(P, N, nch, cumD, cumE, u1, u2):
    """One shot of the alias sampler restated with prefix sums (kernels.cu: k_samp_draw); P(i) reads bin i."""
    dN = float(N)
    idx = min(int(u1 * dN), N - 1)
    f = dN * P(idx)
    q = (N - 1 - idx) // SAMP_CH
    chunk_hi = N - 1 - q * SAMP_CH
    if f < 1.0:
        if u2 < f:
            return idx
        x = 0.0
        for i in range(chunk_hi, idx, -1):
            fi = dN * P(i)
            if fi < 1.0:
                x += 1.0 - fi
        x = cumD[q] + x
        qq = _find_chunk(cumE, nch, x, False)
        while qq < nch:
            hi = N - 1 - qq * SAMP_CH
            e = 0.0
            for i in range(hi, max(hi - SAMP_CH + 1, 0) - 1, -1):
                fi = dN * P(i)
                if fi >= 1.0:
                    e += fi - 1.0
                    if cumE[qq] + e >= x:
                        return i
            qq += 1
        return idx
    e = 0.0
    for i in range(chunk_hi, idx - 1, -1):
        fi = dN * P(i)
        if fi >= 1.0:
            e += fi - 1.0
    Em = cumE[q] + e
    qq = _find_chunk(cumD, nch, Em, True)
    while qq < nch:
        hi = N - 1 - qq * SAMP_CH
        d = 0.0
        for i in range(hi, max(hi - SAMP_CH + 1, 0) - 1, -1):
            fi = dN * P(i)
            if fi < 1.0:
                d += 1.0 - fi
                if cumD[qq] + d > Em:
                    first = 1.0 - (cumD[qq] + d - Em)
                    if u2 >= first:
                        for j in range(idx - 1, -1, -1):
                            if dN * P(j) >= 1.0:
                                return j
                    return idx
        qq += 1
    return idx


def _sampler_worker(rank, world, port, nl, p_slice, seed, shots, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=rank, world_size=world)
        N = world << nl
        nch_l = (1 << nl) // SAMP_CH
        D, E = _chunk_sums(p_slice, N)                          # (1) this rank's chunks, scaled by the WHOLE bin count
        box = [torch.zeros(2 * nch_l, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(box, torch.from_numpy(np.concatenate([D, E])))   # (2) 2 * 2^n / 1024 doubles in total
        gD, gE = np.zeros(world * nch_l), np.zeros(world * nch_l)
        for r in range(world):                                  # global DESCENDING order: the highest rank comes first
            blk = (world - 1 - r) * nch_l
            gD[blk:blk + nch_l] = box[r].numpy()[:nch_l]
            gE[blk:blk + nch_l] = box[r].numpy()[nch_l:]
        cumD, cumE = _exclusive(gD), _exclusive(gE)             # (3) the same scan on every rank
        # (4) every rank draws every shot; bins of other ranks are read from THEIR slices ("peer memory": here a gather)
        sl = [torch.zeros(1 << nl, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(sl, torch.from_numpy(np.ascontiguousarray(p_slice)))
        slices = [t.numpy() for t in sl]
        mask = (1 << nl) - 1
        u = _uniform_doubles(seed, 2 * shots)
        idx = [_draw(lambda i: slices[i >> nl][i & mask], N, world * nch_l, cumD, cumE, u[2 * s], u[2 * s + 1])
               for s in range(shots)]
        dist.destroy_process_group()
        q.put((rank, idx, None))
    except BaseException as ex:  # noqa: BLE001
        q.put((rank, None, repr(ex)))


def test_world2_distributed_sampler_protocol_matches_oracle(oracle_mod):
    """Slices of |psi|^2 on two ranks -> per-slice chunk sums -> all-gather in global descending order -> the same
    scan everywhere -> every rank draws every shot.  The result must be the ORACLE's seeded samples (the true alias
    table built sequentially over all 2^n bins, one mt19937 per call seeded from the device generator)."""
    from dsl import apply_ops, random_circuit
    n, world, shots, dev_seed = 12, 2, 300, 37
    nl = n - 1
    o = oracle_mod.OracleDevice(n, seed=dev_seed)
    apply_ops(o, random_circuit(77, n, 80))
    p = o.probs()
    smp = o.sample(shots)
    want = (smp.astype(np.int64) * (1 << np.arange(n - 1, -1, -1))).sum(axis=1)
    call_seed = int(_mt19937_u32(dev_seed, 1)[0])               # the device generator's first output seeds the call
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sampler_worker, args=(r, world, port, nl, p[r << nl:(r + 1) << nl].copy(), call_seed, shots, q))
             for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
    for pr in procs:
        pr.join(timeout=60)
    for rank, idx, err in res:
        assert err is None, f"rank {rank}: {err}"
        assert np.array_equal(np.asarray(idx, dtype=np.int64), want), f"rank {rank}: samples differ from the oracle's"
