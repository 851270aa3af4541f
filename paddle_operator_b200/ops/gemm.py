"""Fused Linear(+bias)(+ReLU) on tcgen05 tensor cores (csrc/gemm_tcgen05.cu).

    y = linear_act(x, weight, bias, relu=True)        # x [M,K] bf16, weight [N,K] bf16 -> y [M,N] bf16

Forward and the input-gradient GEMM run on the hand-written TMA + tcgen05 kernel
(the ReLU derivative is applied in the dX epilogue of the layer above, so no
separate mask pass exists); the weight-gradient GEMM is a plain library GEMM.
CPU tensors / other dtypes fall back to torch.  Reference parity: no counterpart
in the reference (no compute code); the MLP of deploy/examples/wide_and_deep.yaml.
"""
from __future__ import annotations

from typing import Optional

import os

import torch

from . import _lib
from ._sink import adopt, sink_of


# B operands given as [K, N] (a weight used transposed) are read as stored through an MN-major tcgen05 descriptor;
# PDJ_GEMM_KN=0 makes the callers materialise the transposed copy instead (the A/B arm).
KN = os.environ.get("PDJ_GEMM_KN", "1") == "1"


def linear_bf16(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
                mask: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, b_kn: bool = False) -> torch.Tensor:
    """out[M,N] = epilogue(a[M,K] @ b[N,K]^T), or with `b_kn` epilogue(a[M,K] @ b[K,N]) - e.g. dX = dZ @ W with the
    weight W[out, in] as stored.  a, b: bf16, last dim contiguous, row pitch a multiple of 8."""
    assert a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2 and a.shape[1] == (b.shape[0] if b_kn else b.shape[1])
    if b_kn and not KN:
        b, b_kn = b.t().contiguous(), False
    if a.stride(1) != 1 or a.stride(0) % 8 != 0 or a.data_ptr() % 16 != 0:
        a = _pad_rows(a)
    if b.stride(1) != 1 or b.stride(0) % 8 != 0 or b.data_ptr() % 16 != 0:
        b = _pad_rows(b)
    m, k = a.shape
    n = b.shape[1] if b_kn else b.shape[0]
    if out is None:
        out = torch.empty((m, n), dtype=torch.bfloat16, device=a.device)
    if bias is not None and bias.dtype != torch.float32:
        bias = bias.float()
    if mask is not None:
        assert mask.shape == out.shape and mask.stride() == out.stride() and mask.dtype == torch.bfloat16
    fn = _lib.load().pdj_linear_bf16_kn if b_kn else _lib.load().pdj_linear_bf16
    _lib.check(fn(a.data_ptr(), b.data_ptr(), out.data_ptr(), _lib.ptr(bias), _lib.ptr(mask),
                  m, n, k, a.stride(0), b.stride(0), out.stride(0), int(relu),
                  torch.cuda.current_stream().cuda_stream), "linear_bf16")
    return out


def _pad_rows(t: torch.Tensor) -> torch.Tensor:
    """Copy into a buffer whose row pitch is a multiple of 8 elements (TMA needs 16-byte pitches)."""
    r, c = t.shape
    ld = (c + 7) // 8 * 8
    buf = torch.zeros((r, ld), dtype=t.dtype, device=t.device)
    buf[:, :c].copy_(t)
    return buf[:, :c]


class _LinearActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu, bias_f32):
        y = linear_bf16(x, weight, bias_f32 if bias_f32 is not None else bias, relu=relu)
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.relu = relu
        ctx.has_bias = bias is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        ctx.bias = bias  # the parameter itself: backward only looks at its gradient sink
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        dy = dy.contiguous()
        dz = torch.where(y > 0, dy, torch.zeros((), dtype=dy.dtype, device=dy.device)) if ctx.relu else dy
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dX[M,K] = dZ[M,N] @ W[N,K]  ==  dZ . (W^T)[K,N]^T : the B operand is W^T, K-major over N
            dx = linear_bf16(dz, weight, b_kn=True)
        db_done = False
        need_b = ctx.has_bias and ctx.needs_input_grad[2]
        if ctx.needs_input_grad[1]:
            sink = sink_of(weight)
            dbt = None
            if need_b and ctx.bias_dtype == torch.bfloat16 and dz.stride(1) == 1:  # bias gradient from the same kernel
                bsink = sink_of(ctx.bias)
                dbt = bsink if bsink is not None else torch.empty(dz.shape[1], dtype=torch.bfloat16, device=dz.device)
            res = wgrad_mn(dz, x, out=sink, db_out=dbt)
            dw, db_done = res if dbt is not None else (res, False)
            dw = adopt(dw) if sink is not None else dw
            if db_done:
                db = adopt(dbt) if dbt is sink_of(ctx.bias) else dbt
        if need_b and not db_done:
            from .mlp import colsum

            sink = sink_of(ctx.bias)
            if dz.stride(1) == 1:
                db = colsum(dz, out=sink, dtype=ctx.bias_dtype)
                db = adopt(db) if sink is not None else db
            else:
                db = dz.float().sum(0).to(ctx.bias_dtype)
        return dx, dw, db, None, None


# the bias gradient (column sums of dy) comes out of the weight-gradient kernel itself; PDJ_WGRAD_DB=0 = separate pass
WGRAD_DB = os.environ.get("PDJ_WGRAD_DB", "1") == "1"


def wgrad_mn(dy: torch.Tensor, x: torch.Tensor, out: Optional[torch.Tensor] = None,
             db_out: Optional[torch.Tensor] = None):
    """dW[No, Nc] = dy[M, No]^T @ x[M, Nc] (bf16) on the MN-major tcgen05 kernel (csrc/wgrad_tcgen05.cu: both operands
    read as they sit in memory, accumulators resident in tensor memory, split-K with a fixed-order reduce).  Falls
    back to the library GEMM for shapes the kernel does not take.  `out`: a bf16 tensor whose memory is a dense
    [No, Nc] matrix (e.g. a gradient sink, ops/_sink.py); it is returned.
    `db_out`: a contiguous bf16 [No] tensor that receives the column sums of dy (the bias gradient) from the same
    kernel - one extra N = 16 tensor-core instruction per step against a tile of ones instead of a pass over dy.  With
    `db_out` the result is (dW, filled) where `filled` tells whether the kernel produced it (False: the caller
    computes the bias gradient itself)."""
    m, no = dy.shape
    nc = x.shape[1]
    ok = (dy.is_cuda and dy.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and x.shape[0] == m and no % 8 == 0
          and nc % 8 == 0 and dy.stride(1) == 1 and x.stride(1) == 1 and dy.stride(0) % 8 == 0 and x.stride(0) % 8 == 0
          and dy.data_ptr() % 16 == 0 and x.data_ptr() % 16 == 0 and os.environ.get("PDJ_WGRAD_MN", "1") == "1")
    lib = _lib.load() if ok else None
    with_db = (ok and db_out is not None and WGRAD_DB and db_out.dtype == torch.bfloat16 and db_out.is_contiguous()
               and db_out.numel() == no and db_out.data_ptr() % 8 == 0)
    if with_db:
        nbytes = lib.pdj_wgrad_mn_db_ws_bytes(m, no, nc)
        with_db = nbytes != 0
    if not with_db:
        nbytes = lib.pdj_wgrad_mn_ws_bytes(m, no, nc) if ok else 0
    if nbytes == 0:
        dw = dy.t().mm(x)
        if out is not None:
            out.view(-1).copy_(dw.view(-1)) if out.is_contiguous() else out.copy_(dw.view(out.shape))
            dw = out
        return dw if db_out is None else (dw, False)
    dw = torch.empty((no, nc), dtype=torch.bfloat16, device=x.device) if out is None else out
    assert dw.dtype == torch.bfloat16 and dw.numel() == no * nc and dw.data_ptr() % 16 == 0
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    st = torch.cuda.current_stream().cuda_stream
    if with_db:
        _lib.check(lib.pdj_wgrad_mn_db(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), db_out.data_ptr(), ws.data_ptr(), m, no, nc,
                                       x.stride(0), dy.stride(0), st), "wgrad_mn_db")
    else:
        _lib.check(lib.pdj_wgrad_mn(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), ws.data_ptr(), m, no, nc, x.stride(0),
                                    dy.stride(0), st), "wgrad_mn")
    return dw if db_out is None else (dw, with_db)


def linear_act(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
               bias_f32: Optional[torch.Tensor] = None) -> torch.Tensor:
    if x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16:
        return _LinearActFn.apply(x, weight, bias, relu, bias_f32)
    y = torch.nn.functional.linear(x, weight, bias)
    return torch.relu(y) if relu else y


class FusedLinear(torch.nn.Linear):
    """nn.Linear whose CUDA bf16 forward / dgrad run on the tcgen05 kernel, with an optional fused ReLU."""

    def __init__(self, in_features: int, out_features: int, relu: bool = False, bias: bool = True):
        super().__init__(in_features, out_features, bias=bias)
        self.relu = relu
        self.bias_f32 = None  # a trainer may point this at its fp32 master copy

    def forward(self, x):
        return linear_act(x, self.weight, self.bias, self.relu, self.bias_f32)
