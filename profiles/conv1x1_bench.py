"""Feasibility check: the tcgen05 linear kernel on the 1x1-convolution shapes of
ResNet-50 (batch 256, NHWC bf16) against cuDNN/cuBLAS through F.conv2d, plus the
cost of the separate BN-statistics pass a fused epilogue would remove.

    python profiles/conv1x1_bench.py            # prints one line per shape, JSON at the end
"""
import json
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from paddle_operator_b200.ops import bn as bn_ops  # noqa: E402
from paddle_operator_b200.ops import conv as conv_ops  # noqa: E402
from paddle_operator_b200.ops import gemm  # noqa: E402

B = 256
# (spatial, Cin, Cout, count per step)
SHAPES = [
    (56, 64, 64, 1), (56, 256, 64, 2), (56, 64, 256, 3), (56, 256, 128, 1),
    (28, 512, 128, 3), (28, 128, 512, 4), (28, 512, 256, 1),
    (14, 1024, 256, 5), (14, 256, 1024, 6), (14, 1024, 512, 1),
    (7, 2048, 512, 2), (7, 512, 2048, 3),
]


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    torch.backends.cudnn.benchmark = True
    rows = []
    tot = {"conv": 0.0, "mine": 0.0, "stats": 0.0}
    for hw, cin, cout, cnt in SHAPES:
        x = torch.randn(B, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(cout, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.05).contiguous(memory_format=torch.channels_last)
        x2 = x.permute(0, 2, 3, 1).reshape(-1, cin)
        w2 = w.reshape(cout, cin)
        out = torch.empty(x2.shape[0], cout, device="cuda", dtype=torch.bfloat16)
        y_ref = F.conv2d(x, w)
        y = gemm.linear_bf16(x2, w2, out=out)
        err = (y.float() - y_ref.permute(0, 2, 3, 1).reshape(-1, cout).float()).abs().max().item()
        t_conv = timeit(lambda: F.conv2d(x, w))
        t_mine = timeit(lambda: gemm.linear_bf16(x2, w2, out=out))
        t_mm = timeit(lambda: torch.mm(x2, w2.t(), out=out))
        bn = bn_ops.FusedBnAct(cout, relu=True).cuda()
        yr = y_ref
        t_bn = timeit(lambda: bn(yr))
        y4 = torch.empty_like(y_ref)
        o2 = y4.permute(0, 2, 3, 1).reshape(-1, cout)
        _, part = conv_ops.conv1x1_gemm(x2, w2, stats=True, out=o2)
        err_c = (o2.float() - y_ref.permute(0, 2, 3, 1).reshape(-1, cout).float()).abs().max().item()
        ref_sum = o2.float().sum(0)
        err_s = ((part[:, 0].sum(0) - ref_sum).abs().max() / ref_sum.abs().max().clamp_min(1)).item()
        t_new = timeit(lambda: conv_ops.conv1x1_gemm(x2, w2, stats=True, out=o2))
        t_new_ns = timeit(lambda: conv_ops.conv1x1_gemm(x2, w2, stats=False, out=o2))
        t_bn_pre = timeit(lambda: bn(y4, partial=part))
        mb = (x2.numel() + out.numel()) * 2 / 1e6
        rows.append({"hw": hw, "cin": cin, "cout": cout, "count": cnt, "conv_us": t_conv, "mm_us": t_mm, "mine_us": t_mine,
                     "bn_fwd_us": t_bn, "roof_us": mb / 6.5392, "max_err": err,
                     "new_us": t_new, "new_nostats_us": t_new_ns, "bn_pre_us": t_bn_pre, "err_conv": err_c, "err_sum": err_s})
        tot["stats"] += cnt * (t_new + t_bn_pre)
        tot.setdefault("old_pair", 0.0)
        tot["old_pair"] += cnt * (t_conv + t_bn)
        tot["conv"] += cnt * t_conv
        tot["mine"] += cnt * t_mine
        print(f"{hw:3d} {cin:5d}->{cout:5d} x{cnt}  conv {t_conv:7.1f}  mm {t_mm:7.1f}  mine {t_mine:7.1f}  bn_fwd {t_bn:7.1f}  "
              f"roof {mb / 6.5392:6.1f} us  err {err:.3g} | new {t_new:7.1f} (no stats {t_new_ns:7.1f}) bn_pre {t_bn_pre:7.1f} "
              f"errc {err_c:.3g} errs {err_s:.3g}", flush=True)
    print(json.dumps({"rows": rows, "total_us": tot}))


if __name__ == "__main__":
    main()
