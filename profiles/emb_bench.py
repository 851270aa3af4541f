"""Lock-step timings of the sharded-embedding kernels on N ranks (device events, max over ranks):

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29544 profiles/emb_bench.py

Each kernel is replayed back to back (its own barriers keep the ranks in step), so the number is transfer + kernel,
without the skew that a per-kernel profile of the whole training step folds into the barrier waits.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.parallel.ps import ADAGRAD, ShardedEmbedding  # noqa: E402
from paddle_operator_b200.parallel.symm import Communicator  # noqa: E402
from paddle_operator_b200.workloads.wide_deep import WideDeepJob  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
B, S, D, ROWS = 32768, 26, 16, 1 << 24
comm = Communicator.from_env(WideDeepJob.arena_bytes_needed(ROWS, D, B, world), key=f"embbench-{os.environ.get('MASTER_PORT', '0')}")
emb = ShardedEmbedding(comm, ROWS, D, B, S, has_wide=True, optimizer=ADAGRAD, lr=0.05, nblocks=148)
emb.set_output_layout(432, 13)
g = torch.Generator(device="cuda").manual_seed(rank + 1)
emb.ids_buffer(B).copy_(torch.randint(0, emb.num_rows, (B, S), device="cuda", generator=g))
emb.grad_deep_in(B).normal_(generator=g)
emb.grad_wide_in(B).normal_(generator=g)


def timed(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


res = {}
for how in ("batched", "tma", "loop"):
    os.environ["PDJ_EMB_LOOKUP"] = how
    res[f"lookup[{how}]"] = timed(lambda: emb.lookup_raw(B))
os.environ.pop("PDJ_EMB_LOOKUP")
emb.async_sgd = True  # push only (mode 1 adds into the weights; same traffic as staging)
for how in ("batched", "loop"):
    os.environ["PDJ_EMB_PUSH"] = how
    res[f"push[{how}]"] = timed(lambda: emb.push_raw(B, 1e-6))
os.environ.pop("PDJ_EMB_PUSH")
emb.async_sgd = False
res["push+apply"] = timed(lambda: emb.push_raw(B, 1e-6))
res["barrier"] = timed(lambda: comm.barrier())
if rank == 0:
    remote = B * S * (world - 1) / world
    print(f"world={world}  rows/rank={B * S}  remote rows/rank={int(remote)}  ({remote * 68 / 1e6:.1f} MB pulled, {remote * 85 / 1e6:.1f} MB pushed)")
    for k, v in res.items():
        print(f"  {k:16s} {v:8.1f} us")
comm.close()
dist.destroy_process_group()
