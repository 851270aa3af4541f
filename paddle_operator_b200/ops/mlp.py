"""A ReLU MLP whose forward and backward chain the tcgen05 fused-linear kernel.

    h_0 = x;  h_l = relu(h_{l-1} W_l^T + b_l)          l = 1..L

Forward: L launches of the tcgen05 kernel (bias + ReLU in the epilogue).
Backward, per layer from the top:
    db_l = colsum(dz_l)                                 (csrc/optim_misc.cu)
    dW_l = dz_l^T h_{l-1}                               MN-major tcgen05 kernel (csrc/wgrad_tcgen05.cu)
    dz_{l-1} = (dz_l W_l) * (h_{l-1} > 0)               tcgen05 kernel, ReLU derivative as epilogue mask
so the only stand-alone mask pass is the one on the incoming gradient.

Reference parity: no counterpart (the reference has no compute code); the deep tower of the
Wide&Deep workload, deploy/examples/wide_and_deep.yaml.
"""
from __future__ import annotations

import torch

from . import _lib
from .gemm import linear_bf16, wgrad_mn


def colsum(x: torch.Tensor) -> torch.Tensor:
    """fp32 column sums of a 2-D bf16 CUDA tensor (last dim contiguous)."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    out = torch.empty(x.shape[1], dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().pdj_colsum_bf16(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], x.stride(0),
                                           torch.cuda.current_stream().cuda_stream), "colsum_bf16")
    return out


class _FusedMLPFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, n_layers, top_grad_masked, *params):
        ws = params[:n_layers]
        bs = params[n_layers:2 * n_layers]
        b32 = params[2 * n_layers:]
        hs = [x]
        for l in range(n_layers):
            bias = b32[l] if b32[l] is not None else bs[l]
            hs.append(linear_bf16(hs[-1], ws[l], bias, relu=True))
        ctx.save_for_backward(*hs, *ws)
        ctx.n = n_layers
        ctx.top_grad_masked = top_grad_masked
        ctx.bias_dtypes = [b.dtype for b in bs]
        return hs[-1]

    @staticmethod
    def backward(ctx, dy):
        n = ctx.n
        saved = ctx.saved_tensors
        hs, ws = saved[:n + 1], saved[n + 1:]
        if ctx.top_grad_masked:  # the consumer already zeroed dy where the top activation is zero (ops.wd_head)
            dz = dy.contiguous()
        else:
            dz = torch.where(hs[n] > 0, dy, torch.zeros((), dtype=dy.dtype, device=dy.device)).contiguous()
        dws, dbs = [None] * n, [None] * n
        dx = None
        for l in range(n, 0, -1):  # layer l maps hs[l-1] -> hs[l] with ws[l-1]
            dbs[l - 1] = colsum(dz).to(ctx.bias_dtypes[l - 1])
            dws[l - 1] = wgrad_mn(dz, hs[l - 1])
            if l > 1:
                dz = linear_bf16(dz, ws[l - 1].t().contiguous(), mask=hs[l - 1])
            elif ctx.needs_input_grad[0]:
                dx = linear_bf16(dz, ws[0].t().contiguous())
        return (dx, None, None, *dws, *dbs, *([None] * n))


class FusedMLP(torch.nn.Module):
    """Stack of Linear+ReLU layers.  CUDA bf16 inputs run on the hand-written kernels; anything
    else (CPU tests, fp32) uses torch ops with identical parameters."""

    def __init__(self, dims):
        super().__init__()
        self.layers = torch.nn.ModuleList([torch.nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:])])
        self.bias_f32 = [None] * len(self.layers)  # a trainer may point these at fp32 master copies

    def forward(self, x, top_grad_masked: bool = False):
        """`top_grad_masked`: the gradient that will arrive for the output is already zero wherever the output
        is zero (a fused consumer applied the ReLU derivative), so backward skips its own mask pass."""
        if x.is_cuda and x.dtype == torch.bfloat16 and self.layers[0].weight.dtype == torch.bfloat16:
            ws = [l.weight for l in self.layers]
            bs = [l.bias for l in self.layers]
            return _FusedMLPFn.apply(x, len(ws), bool(top_grad_masked), *ws, *bs, *self.bias_f32)
        for l in self.layers:
            x = torch.relu(l(x))
        return x
