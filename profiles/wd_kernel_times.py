"""Per-kernel device times of the Wide&Deep step from torch.profiler (CUPTI), any number of ranks:

    python profiles/wd_kernel_times.py                                   # one GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 profiles/wd_kernel_times.py

One GPU: eager steps; several ranks: graph replays (WD_GRAPH=1, eager ranks drift apart and the barrier waits would
dominate).  Warm caches, real NVLink traffic - the view ncu cannot
give for a multi-rank run.  Rank 0 prints the table (us per step, averaged over the profiled steps).
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.workloads.wide_deep import WideDeepJob  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    import torch.distributed as dist

    dist.init_process_group("nccl")
    job = WideDeepJob.from_env(batch=32768, use_graph=os.environ.get("WD_GRAPH", "1") == "1")
else:
    job = WideDeepJob(batch=32768, use_graph=False)
job.prepare()
for _ in range(5):
    job.step_device()
torch.cuda.synchronize()
STEPS = 10
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    for _ in range(STEPS):
        job.step_device()
    torch.cuda.synchronize()
if rank == 0:
    rows = [(e.key, e.device_time_total / STEPS, e.count / STEPS) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"world={world}: {tot:.1f} us of kernels per step")
    for k, t, n in rows[:24]:
        print(f"{t:8.1f} us  x{n:4.1f}  {k[:100]}")
if world > 1:
    dist.barrier()
    job.close() if hasattr(job, "close") else None
    dist.destroy_process_group()
