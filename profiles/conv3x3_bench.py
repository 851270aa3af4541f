"""Times the tcgen05 3x3-convolution kernels against cuDNN on the ResNet-50 layer shapes at batch 256
(CUDA events, L2 flushed between iterations by a 256 MB write).

    python profiles/conv3x3_bench.py [--batch 256]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from paddle_operator_b200.ops import conv as C  # noqa: E402


def timeit(fn, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    args = ap.parse_args()
    cl = torch.channels_last
    rows = []
    for hw, c in [(56, 64), (28, 128), (14, 256), (7, 512)]:
        x = torch.randn(args.batch, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
        dy = torch.randn(args.batch, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
        w = (torch.randn(c, c, 3, 3, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=cl)
        flops = 2.0 * args.batch * hw * hw * c * c * 9
        row = {"hw": hw, "c": c, "gflop": round(flops / 1e9, 1)}
        row["wgrad_ours_us"] = round(timeit(lambda: C.conv3x3_wgrad(x, dy, w.shape)), 1)
        row["wgrad_cudnn_us"] = round(timeit(lambda: torch.ops.aten.convolution_backward(
            dy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [False, True, False])), 1)
        row["wgrad_ours_tflops"] = round(flops / row["wgrad_ours_us"] / 1e6, 1)
        if hasattr(C, "conv3x3_fprop"):
            row["fprop_ours_us"] = round(timeit(lambda: C.conv3x3_fprop(x, w)), 1)
            row["dgrad_ours_us"] = round(timeit(lambda: C.conv3x3_dgrad(dy, w)), 1)
        row["fprop_cudnn_us"] = round(timeit(lambda: torch.nn.functional.conv2d(x, w, padding=1)), 1)
        row["dgrad_cudnn_us"] = round(timeit(lambda: torch.ops.aten.convolution_backward(
            dy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [True, False, False])), 1)
        rows.append(row)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
