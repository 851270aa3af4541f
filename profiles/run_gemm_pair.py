"""One 8192 x 4096 x 4096 fused linear in each mode (one CTA per tile, CTA pairs) for ncu."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.ops.gemm import linear_bf16  # noqa: E402

a = torch.randn(8192, 4096, device="cuda").to(torch.bfloat16)
w = torch.randn(4096, 4096, device="cuda").to(torch.bfloat16)
bias = torch.randn(4096, device="cuda")
for pair in ("0", "1", "0", "1"):
    os.environ["PDJ_LINEAR_PAIR"] = pair
    linear_bf16(a, w, bias, relu=True)
torch.cuda.synchronize()
print("ok")
