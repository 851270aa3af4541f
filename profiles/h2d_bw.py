"""Pinned host -> device copy rate of this box (what bounds the end-to-end Wide&Deep number)."""
import torch

for mb in (1, 8, 64):
    h = torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    d = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print(f"H2D {mb} MiB: {mb * 20 / 1024 / (e0.elapsed_time(e1) / 1e3):.1f} GiB/s")
