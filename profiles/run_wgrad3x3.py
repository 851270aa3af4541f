"""ncu driver: one 3x3 weight-gradient launch per ResNet-50 layer shape at batch 256 (plus a warm-up each)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.ops import conv as C  # noqa: E402

cl = torch.channels_last
for hw, c in [(56, 64), (28, 128), (14, 256), (7, 512)]:
    x = torch.randn(256, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    dy = torch.randn(256, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    for _ in range(2):
        C.conv3x3_wgrad(x, dy, (c, c, 3, 3))
    torch.cuda.synchronize()
print("ok")
