"""Under torchrun: time the pieces of the sharded end-to-end step separately (circuit, Expval, PartialProbs)."""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch.distributed as dist
from catalyst_b200 import capi, workloads
from dsl import apply_ops
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dist.init_process_group("gloo")
box = [capi.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 10
ops = workloads.c5_random_depth(n, depth=depth)
d = capi.B200Device(n, seed=37, world_size=world, rank=rank, comm="nccl:" + box[0], device=lr)
for rep in range(5):
    dist.barrier()
    t0 = time.perf_counter()
    d.set_basis_state([0] * n, list(range(n)))
    apply_ops(d, ops)
    t1 = time.perf_counter()
    if rep % 2 == 0:
        d.synchronize()
    t2 = time.perf_counter()
    ez = d.expval(d.obs_named("PauliZ", 0)); t3 = time.perf_counter()
    pp = d.probs(list(range(10))); t4 = time.perf_counter()
    s = d.stats()
    if rank == 0:
        print("rep %d (%s) enqueue %.0f ms, sync %.0f ms, expval %.0f ms, probs %.0f ms, total %.0f ms; exchanges %d" % (
            rep, "sync before expval" if rep % 2 == 0 else "no sync", 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2),
            1e3 * (t4 - t3), 1e3 * (t4 - t0), s.exchanges), flush=True)
    d.jit_wait()
d.close()
dist.destroy_process_group()
