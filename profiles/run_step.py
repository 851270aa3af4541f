"""Runs ResNet-50 training steps eagerly (no CUDA graph) for ncu.  The profiled
window is the step after warm-up, bracketed by cudaProfilerStart/Stop, so use
`ncu --profile-from-start off ...`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.workloads.resnet import ResNetJob  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.cuda.set_device(0)
job = ResNetJob(batch=batch, use_graph=False)
job.prepare()
for _ in range(3):
    job.step_device()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
for _ in range(steps):
    job.step_device()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("loss", float(job.trainer.loss.item()))
