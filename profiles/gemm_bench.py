"""Times the tcgen05 fused linear kernel against torch (cuBLAS + separate bias/ReLU kernels)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.ops.gemm import linear_bf16  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    t = 0.0
    for _ in range(n):
        if os.environ.get("NOFLUSH") is None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t += e0.elapsed_time(e1)
    return t / n * 1e3  # us


for (m, n, k) in [(32768, 400, 432), (32768, 400, 400), (32768, 1, 400), (8192, 4096, 4096), (16384, 1000, 2048)]:
    a = torch.randn(m, (k + 7) // 8 * 8, device=dev).to(torch.bfloat16)[:, :k]
    w = torch.randn(n, (k + 7) // 8 * 8, device=dev).to(torch.bfloat16)[:, :k]
    wc = w.contiguous()
    ac = a.contiguous()
    bias = torch.randn(n, device=dev)
    bias16 = bias.to(torch.bfloat16)
    os.environ["PDJ_LINEAR_PAIR"] = "0"
    t_ours = timeit(lambda: linear_bf16(a, w, bias, relu=True))
    os.environ["PDJ_LINEAR_PAIR"] = "1"  # CTA pairs (tcgen05 cta_group::2); read by the library on every call
    t_pair = timeit(lambda: linear_bf16(a, w, bias, relu=True))
    os.environ.pop("PDJ_LINEAR_PAIR")
    t_torch = timeit(lambda: torch.relu(torch.nn.functional.linear(ac, wc, bias16)))
    t_mm = timeit(lambda: torch.mm(ac, wc.t()))
    fl = 2.0 * m * n * k
    print(f"M={m} N={n} K={k}: ours(fused) {t_ours:8.1f} us {fl / t_ours / 1e6:7.1f} TFLOP/s | CTA pairs {t_pair:8.1f} us {fl / t_pair / 1e6:7.1f} TFLOP/s | torch linear+relu {t_torch:8.1f} us "
          f"| cuBLAS mm only {t_mm:8.1f} us {fl / t_mm / 1e6:7.1f} TFLOP/s", flush=True)
