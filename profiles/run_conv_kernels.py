"""One launch each of the tcgen05 convolution kernels and the TMA-streamed BN reductions on ResNet-50
layer shapes (batch 256), bracketed by cudaProfilerStart/Stop for `ncu --profile-from-start off --set full`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.ops import bn as bn_ops  # noqa: E402
from paddle_operator_b200.ops import conv as conv_ops  # noqa: E402

cl = torch.channels_last
B = 256
dev = "cuda"


def t(n, c, h, w, scale=1.0):
    return (torch.randn(n, c, h, w, device=dev) * scale).to(torch.bfloat16).contiguous(memory_format=cl)


x64 = t(B, 64, 56, 56)
w64_256 = t(256, 64, 1, 1, 0.05)
dy64 = t(B, 64, 56, 56)
skip256 = t(B, 256, 56, 56)
w256_64 = t(64, 256, 1, 1, 0.05)
xs = t(B, 16, 115, 115)
ws = t(64, 16, 4, 4, 0.05)
dys = t(B, 64, 112, 112)
bn = bn_ops.FusedBnAct(256, relu=True).to(dev)
x256 = t(B, 256, 56, 56).requires_grad_(True)
res = t(B, 256, 56, 56)


def run():
    # 1x1 conv 64 -> 256 with BN partial sums (MODE_STATS)
    y, partial = conv_ops.conv1x1_gemm(conv_ops._rows(x64), w64_256.reshape(256, 64), stats=True)
    # conv1 dgrad + skip-gradient add: dX[M,256] = skip + dY[M,64] @ W[64,256]  (MODE_ADD)
    conv_ops.conv1x1_gemm_add(conv_ops._rows(dy64), w256_64.reshape(64, 256).t().contiguous(), conv_ops._rows(skip256))
    # stem forward + statistics, stem weight gradient
    xa, wa = xs.detach().requires_grad_(False), ws.detach().requires_grad_(True)
    ys, ps = conv_ops.stem_s2d_conv(xa, wa)
    ys.backward(dys)
    # BN forward (TMA statistics + apply) and backward (TMA reduce + apply) on the widest layer-1 tensor
    z = bn(x256, res)
    z.backward(skip256)


for _ in range(2):
    run()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
run()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok")
