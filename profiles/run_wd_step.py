"""Runs Wide&Deep steps eagerly for ncu (profile window = one step after warm-up)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.workloads.wide_deep import WideDeepJob  # noqa: E402

torch.cuda.set_device(0)
job = WideDeepJob(batch=32768, use_graph=False, rows=int(os.environ.get("WD_ROWS", str(1 << 24))))
job.prepare()
for _ in range(3):
    job.step_device()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
job.step_device()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("loss", float(job.trainer.loss.item()))
