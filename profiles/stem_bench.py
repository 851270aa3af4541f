"""Micro-benchmark: the ResNet stem convolution (7x7/2, 64 filters) forward +
backward for input channels padded from 3 to 4 or 8 (NHWC bf16, batch 256)."""
import torch

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda", 0)
for cin in (3, 4, 8):
    conv = torch.nn.Conv2d(cin, 64, 7, stride=2, padding=3, bias=False).to(dev).to(torch.bfloat16).to(
        memory_format=torch.channels_last)
    x = torch.randn(256, cin, 224, 224, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    for need_dx in (False,):
        x.requires_grad_(need_dx)
        for _ in range(5):
            y = conv(x)
            y.backward(y)
        torch.cuda.synchronize()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        n = 20
        tf = tb = 0.0
        for _ in range(n):
            e0.record()
            y = conv(x)
            e1.record()
            y.backward(y)
            e2.record()
            torch.cuda.synchronize()
            tf += e0.elapsed_time(e1)
            tb += e1.elapsed_time(e2)
        print(f"cin={cin} fwd {tf / n:.3f} ms  bwd(wgrad) {tb / n:.3f} ms", flush=True)
