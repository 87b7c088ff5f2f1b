#!/usr/bin/env python
"""A Python model of the host-side planner (catalyst_b200/csrc/simulator.cpp: `run_ops` and the sharded `execute`),
for exploring planning heuristics without a GPU and without compiling anything.

It reproduces the C++ planner's numbers: C3 at 30 qubits 17 passes (greedy) and C5 47, and for the sharded 34-qubit C5
of the benchmark 69 passes + 7 three-bit swaps on 8 ranks, 68 + 6 on 4, 62 + 5 on 2 (the values measured in
profiles/r2_v3_bench_n*.json, `exchanges_per_step`).  tests/test_planner_dryrun.py keeps model and C++ in step.

    python scripts/planner_model.py            # the table of DESIGN.md section 5 ("planner model")
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

T = 11          # tile bits
CAP = 160       # ops per pass (Simulator::pass_op_cap_)


def popc(x):
    return bin(x).count("1")


def fit_rows(req, t=T):
    """Largest c such that the bits of req at or above c fit the t - c scattered tile bits (simulator.cpp: fit_rows)."""
    if popc(req) > t:
        return -1
    c = t
    while c > 0 and popc(req >> c) > t - c:
        c -= 1
    return c


def cost(c):
    """Relative duration of a pass by row width 2^c amplitudes (the planner's measured table)."""
    return 1.0 if c >= 7 else {6: 1.12, 5: 1.2, 4: 1.3, 3: 1.38, 2: 1.6}.get(c, 3.0)


def canon(ops, n):
    """(xmask, zsupport) per canonical op, in bit space; uncontrolled SWAPs relabel wires; adjacent one-qubit
    non-diagonal ops on the same bit merge (Simulator::enqueue)."""
    wb = list(range(n - 1, -1, -1))
    out = []
    for name, _, wires in ops:
        b = [wb[w] for w in wires]
        if name in ("RX", "RY", "Hadamard", "PauliX", "PauliY"):
            out.append((1 << b[0], 0))
        elif name in ("RZ", "T", "S", "PauliZ", "PhaseShift"):
            out.append((0, 1 << b[0]))
        elif name == "CNOT":
            out.append((1 << b[1], 1 << b[0]))
        elif name in ("CZ", "ControlledPhaseShift"):
            out.append((0, (1 << b[0]) | (1 << b[1])))
        elif name == "SWAP":
            wb[wires[0]], wb[wires[1]] = wb[wires[1]], wb[wires[0]]
        else:
            raise ValueError(name)
    merged = []
    for op in out:
        if merged and op[1] == 0 and merged[-1][1] == 0 and op[0] == merged[-1][0] and op[0]:
            continue
        merged.append(op)
    return merged


def plan(ops, cmin=2, iters=6, width=3):
    """Passes [(number of ops, row bits c)] for one queue.  iters = 0: the greedy planner with adaptive rows;
    width = 1: refinement over any bit (plan_refine_mode 1); width >= 2: over scattered bits, `width` candidates."""
    done = [False] * len(ops)
    first, remaining, passes = 0, len(ops), []

    def build(cm, forbid=0):
        req = bx = bz = 0
        idx, c_cur, use = [], T, {}
        for i in range(first, min(len(ops), first + 4096)):
            if len(idx) >= CAP:
                break
            if done[i]:
                continue
            xs, zs = ops[i][0], ops[i][1]
            if not (xs & forbid) and not (xs & (bx | bz)) and not (zs & bx):
                nreq = req | xs
                c = fit_rows(nreq)
                if c >= 0 and (c >= cm or not idx):
                    req, c_cur = nreq, c
                    idx.append(i)
                    m = xs
                    while m:
                        b = m & -m
                        use[b] = use.get(b, 0) + 1
                        m &= m - 1
                    continue
            bx |= xs
            bz |= zs
        return idx, req, c_cur, use, cm

    def score(cand):
        return len(cand[0]) / cost(cand[2])

    while remaining:
        best = None
        for cm in range(cmin, 8):
            cand = build(cm)
            if best is None or score(cand) > score(best) * 1.0001:
                best = cand
        cur, forbid = best, 0
        for _ in range(iters):
            keep = ops[cur[0][0]][0] if cur[0] else 0
            v = sorted((u, b) for b, u in cur[3].items() if not (b & keep) and (width == 1 or b >= (1 << cur[2])))
            if not v:
                break
            trials = [(build(cur[4], forbid | b), b) for _, b in v[:width]]
            nxt, b = max(trials, key=lambda t: score(t[0]))
            forbid |= b
            if score(nxt) > score(best) * 1.0001:
                best = nxt
            cur = nxt
        for i in best[0]:
            done[i] = True
        remaining -= len(best[0])
        while first < len(ops) and done[first]:
            first += 1
        passes.append((len(best[0]), best[2]))
    return passes


def to_phys(pos, m):
    r = 0
    while m:
        q = (m & -m).bit_length() - 1
        m &= m - 1
        r |= 1 << pos[q]
    return r


def execute_sharded(ops, n, g, **kw):
    """The sharded executor: segments of ops runnable under the current layout, then the swap that makes the first
    blocked op runnable (all rank-bit qubits with a pending non-diagonal use come in, evicting the local qubits whose
    next non-diagonal use is farthest).  Returns (passes, [bits per swap])."""
    nl = n - g
    pos = list(range(n))
    done = [False] * len(ops)
    first, remaining, passes, swaps = 0, len(ops), [], []
    local = (1 << nl) - 1
    while remaining:
        bx = bz = 0
        seg = []
        for i in range(first, len(ops)):
            if done[i]:
                continue
            xs, zs = ops[i]
            if not (xs & (bx | bz)) and not (zs & bx) and not (to_phys(pos, xs) & ~local):
                seg.append((to_phys(pos, xs), to_phys(pos, zs)))
                done[i] = True
                remaining -= 1
                continue
            bx |= xs
            bz |= zs
        while first < len(ops) and done[first]:
            first += 1
        if seg:
            passes += plan(seg, **kw)
        if not remaining:
            break
        head = ops[first]
        INF = 10 ** 9
        nu = [INF] * n
        for i in range(first, len(ops)):
            if done[i]:
                continue
            m = ops[i][0]
            while m:
                q = (m & -m).bit_length() - 1
                m &= m - 1
                if nu[q] == INF:
                    nu[q] = i
        inq = sorted((q for q in range(n) if pos[q] >= nl and nu[q] < INF), key=lambda q: nu[q])
        victims = sorted((q for q in range(n) if pos[q] < nl and not (head[0] >> q) & 1), key=lambda q: (-nu[q], -pos[q]))
        pairs, vi = [], 0
        for q in inq:
            if vi >= len(victims):
                break
            v = victims[vi]
            if not ((head[0] >> q) & 1) and nu[v] <= nu[q]:
                continue
            pairs.append((q, v))
            vi += 1
        assert pairs
        for q, v in pairs:
            pos[q], pos[v] = pos[v], pos[q]
        swaps.append(len(pairs))
    return passes, swaps


def main():
    from catalyst_b200 import workloads
    variants = (("round-2 bench runs (rows >= 128 B, greedy)", dict(cmin=3, iters=0)),
                ("64-byte rows", dict(cmin=2, iters=0)),
                ("+ refinement, any bit (mode 1)", dict(cmin=2, iters=6, width=1)),
                ("+ refinement, scattered bits x3 (mode 2, HEAD)", dict(cmin=2, iters=6, width=3)))
    for tag, n, ops in (("C3", 30, workloads.c3_qft_random(30)), ("C5", 30, workloads.c5_random_depth(30, depth=40))):
        co = canon(ops, n)
        for label, kw in variants:
            p = plan(co, **kw)
            print(f"{tag} n={n} one GPU   {label:48s} passes {len(p):3d}  cost-weighted {sum(cost(c) for _, c in p):5.1f}")
    co = canon(workloads.c5_random_depth(34, depth=40), 34)
    for g in (1, 2, 3):
        for label, kw in variants:
            p, s = execute_sharded(co, 34, g, **kw)
            print(f"C5 n=34 on {1 << g} GPUs {label:48s} passes {len(p):3d}  cost-weighted {sum(cost(c) for _, c in p):5.1f}"
                  f"  swaps {len(s)} x {s[0]} bit(s)")


if __name__ == "__main__":
    main()
