"""Max pooling for channels-last bf16 (csrc/pool.cu): forward saves a one-byte
arg-max per element, backward is a deterministic gather.  CPU / other dtypes use
torch.  Reference parity: no counterpart in the reference (no model code)."""
from __future__ import annotations

import torch

from . import _lib


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p):
        if not x.is_contiguous(memory_format=torch.channels_last):
            x = x.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        ho, wo = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
        y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device)
        _lib.check(_lib.load().pdj_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), n, h, w, c, k, s, p,
                                               torch.cuda.current_stream().cuda_stream), "maxpool_fwd")
        ctx.save_for_backward(idx)
        ctx.geom = (n, c, h, w, k, s, p)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, h, w, k, s, p = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((n, c, h, w), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        _lib.check(_lib.load().pdj_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), n, h, w, c, k, s, p,
                                               torch.cuda.current_stream().cuda_stream), "maxpool_bwd")
        return dx, None, None, None


def max_pool2d_nhwc(x: torch.Tensor, kernel_size: int = 3, stride: int = 2, padding: int = 1) -> torch.Tensor:
    if x.is_cuda and x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0:
        return _MaxPoolFn.apply(x, kernel_size, stride, padding)
    return torch.nn.functional.max_pool2d(x, kernel_size, stride, padding)


class FusedMaxPool2d(torch.nn.Module):
    def __init__(self, kernel_size: int = 3, stride: int = 2, padding: int = 1):
        super().__init__()
        self.k, self.s, self.p = kernel_size, stride, padding

    def forward(self, x):
        return max_pool2d_nhwc(x, self.k, self.s, self.p)
