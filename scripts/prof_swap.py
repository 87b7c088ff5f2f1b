"""Micro-benchmark of the global<->local qubit swap and of a tile pass on a sharded register, one process,
one thread per GPU: `python scripts/prof_swap.py [n] [world] [mode]`, mode = p2p (in-place pairwise kernel over mapped
shards), push (one kernel writes into every peer's staging) or staged (pack -> device-to-device copies -> unpack)."""
import sys, time, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from catalyst_b200 import capi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mode = sys.argv[3] if len(sys.argv) > 3 else "p2p"
kw = {"p2p": dict(p2p_swap=1, swap_push=0), "push": dict(p2p_swap=0, swap_push=1), "staged": dict(p2p_swap=0, swap_push=0)}[mode]
g = world.bit_length() - 1


def prog(d, r):
    d.allocate_qubits(n)
    out = []
    def timed(label, fn, reps=3):
        for _ in range(reps):
            d.synchronize(); d.reset_stats()
            t0 = time.perf_counter(); fn(); d.synchronize(); dt = time.perf_counter() - t0
            s = d.stats()
        out.append("%s rank %d %-34s %.1f ms  passes %d exchanges %d (in-place %d, push %d) sent %.1f GiB" % (
            mode, r, label, 1e3 * dt, s.tile_passes, s.exchanges, s.p2p_exchanges, s.swap_push_exchanges, s.exchange_bytes / 2**30))
    timed("local H (1 pass)", lambda: d.named_op("Hadamard", [], [n - 1]))
    timed("H on a middle local wire", lambda: d.named_op("Hadamard", [], [g + 1]))
    # a Hadamard on every wire needs every qubit local at some point: >= 1 swap per repetition
    def allh():
        for w in range(n):
            d.named_op("Hadamard", [], [w])
    for rep in range(3):
        timed("H on all %d wires, rep %d" % (n, rep), allh, reps=1)
    return out


for lines in capi.run_spmd(world, prog, devices=list(range(world)), jit="off", **kw):
    print("\n".join(lines))
