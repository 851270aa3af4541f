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

import ctypes
from typing import Optional

import torch

from . import _lib
from . import _sink
from ._sink import adopt, sink_of
from .gemm import linear_bf16, wgrad_mn


_CS_WS: dict = {}


def _cs_workspace(device, nbytes: int) -> torch.Tensor:
    """Persistent partial-sum rows + ticket counter of the column-sum kernel (zeroed once; the kernel leaves the
    ticket at zero).  Created on the first eager call, i.e. before any graph capture."""
    key = (device.type, device.index)
    ws = _CS_WS.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _CS_WS[key] = ws
    return ws


def colsum_w(x: torch.Tensor, dests, weights: Optional[torch.Tensor] = None, gscale: Optional[torch.Tensor] = None,
             extra: Optional[torch.Tensor] = None, ones: bool = False) -> None:
    """Deterministic weighted column sums in one launch (csrc/optim_misc.cu, colsum_w_kernel):
    res = [ w^T x | w^T extra | sum(w) ],  w = weights * gscale (each optional, default 1).
    `dests`: up to four (tensor, first result column) pairs; each contiguous bf16 / fp32 tensor receives
    tensor.numel() result columns."""
    lib = _lib.load()
    m, n = x.shape
    ne = 0 if extra is None else extra.shape[1]
    assert x.is_cuda and x.dtype == torch.bfloat16 and x.stride(1) == 1 and 1 <= len(dests) <= 4
    assert weights is None or (weights.dtype == torch.float32 and weights.is_contiguous() and weights.numel() == m)
    assert extra is None or (extra.dtype == torch.float32 and extra.stride(1) == 1 and extra.shape[0] == m)
    ws = _cs_workspace(x.device, lib.pdj_colsum_w_ws_bytes(m, n, ne, int(ones)))
    k = len(dests)
    for t, _ in dests:
        assert t.is_contiguous() and t.dtype in (torch.bfloat16, torch.float32)
    ptrs = (ctypes.c_void_p * k)(*[t.data_ptr() for t, _ in dests])
    starts = (ctypes.c_int * k)(*[int(c0) for _, c0 in dests])
    counts = (ctypes.c_int * k)(*[t.numel() for t, _ in dests])
    bf = (ctypes.c_int * k)(*[int(t.dtype == torch.bfloat16) for t, _ in dests])
    _lib.check(lib.pdj_colsum_w(x.data_ptr(), x.stride(0), m, n, _lib.ptr(weights), _lib.ptr(gscale), _lib.ptr(extra),
                                0 if extra is None else extra.stride(0), ne, int(ones), ws.data_ptr(), ptrs, starts,
                                counts, bf, k, torch.cuda.current_stream().cuda_stream), "colsum_w")


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None, dtype=torch.float32) -> torch.Tensor:
    """Column sums of a 2-D bf16 CUDA tensor (last dim contiguous) into `out` (bf16 or fp32, default a new `dtype`
    tensor)."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    if out is None:
        out = torch.empty(x.shape[1], dtype=dtype, device=x.device)
    if x.shape[1] % 8 == 0 and x.stride(0) % 8 == 0 and x.data_ptr() % 16 == 0 and x.shape[1] <= 1000:
        colsum_w(x, [(out, 0)])
        return out
    tmp = out if out.dtype == torch.float32 else torch.empty(x.shape[1], dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().pdj_colsum_bf16(x.data_ptr(), tmp.data_ptr(), x.shape[0], x.shape[1], x.stride(0),
                                           torch.cuda.current_stream().cuda_stream), "colsum_bf16")
    if tmp is not out:
        out.copy_(tmp)
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
        ctx.biases = list(bs)  # parameters (not saved tensors): only their gradient sinks are looked at
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
        # input-gradient chain first: dz_{l-1} = (dz_l W_l) * (h_{l-1} > 0), the weight read as stored
        dzs = [None] * (n + 1)
        dzs[n] = dz
        for l in range(n, 0, -1):  # layer l maps hs[l-1] -> hs[l] with ws[l-1]
            if l > 1:
                dzs[l - 1] = linear_bf16(dzs[l], ws[l - 1], mask=hs[l - 1], b_kn=True)
            elif ctx.needs_input_grad[0]:
                dx = linear_bf16(dzs[1], ws[0], b_kn=True)

        def param_grads():
            for l in range(n, 0, -1):
                # both parameter gradients land in the trainer's flat gradient buffer when it provides sinks
                bsink, wsink = sink_of(ctx.biases[l - 1]), sink_of(ws[l - 1])
                need_b, need_w = ctx.needs_input_grad[3 + n + l - 1], ctx.needs_input_grad[3 + l - 1]
                db_done = False
                if need_w:
                    # a bf16 bias gradient comes out of the weight-gradient kernel (column sums of dz against a tile
                    # of ones on the tensor cores) - no separate pass over dz
                    db = None
                    if need_b and ctx.bias_dtypes[l - 1] == torch.bfloat16:
                        db = bsink if bsink is not None else torch.empty(dzs[l].shape[1], dtype=torch.bfloat16,
                                                                         device=dzs[l].device)
                    res = wgrad_mn(dzs[l], hs[l - 1], out=wsink, db_out=db)
                    dw, db_done = res if db is not None else (res, False)
                    dws[l - 1] = adopt(dw) if wsink is not None else dw
                    if db_done:
                        dbs[l - 1] = adopt(db) if bsink is not None else db
                if need_b and not db_done:
                    db = colsum(dzs[l], out=bsink, dtype=ctx.bias_dtypes[l - 1])
                    dbs[l - 1] = adopt(db) if bsink is not None else db

        if _sink.overlap_enabled():
            # the parameter gradients run on a side stream while the main stream continues with whatever consumes dx
            # (the embedding push / apply); the trainer waits for the event before it touches the gradients
            main, side = torch.cuda.current_stream(), _sink.side_stream(dy.device)
            fork = torch.cuda.Event()
            fork.record(main)
            side.wait_event(fork)
            with torch.cuda.stream(side):
                param_grads()
                done = torch.cuda.Event()
                done.record(side)
            _sink.register_side_work(done, *dzs, *hs)
        else:
            param_grads()
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
