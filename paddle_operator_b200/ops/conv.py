"""1x1 convolution on the tcgen05 GEMM pipeline, fused with BatchNorm statistics.

A stride-1 1x1 convolution over a channels-last activation is the GEMM
[N*H*W, Cin] x [Cout, Cin]^T.  `conv1x1_stats` runs it on the hand-written
TMA + tcgen05 kernel (csrc/conv_tcgen05.cu::conv1x1_tcgen05_kernel) whose
epilogue also accumulates the per-channel sum and sum of squares of the output,
so the training-mode BatchNorm that follows (ops.bn, `partial=`) skips its
statistics pass - one full read of the activation per layer.

    y, partial = conv1x1_stats(x, weight)          # x NHWC bf16, weight [Cout, Cin, 1, 1] bf16
    z = bn(y, residual, partial=partial)           # FusedBnAct

Backward: input and weight gradients are plain library convolutions
(aten.convolution_backward), exactly what autograd would have run.

Reference parity: none - the reference launches a user image that runs the
ResNet of deploy/examples/resnet.yaml:13-19; this is part of the compute layer
the runtime owns.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import _lib
from ._sink import adopt, sink_of


# The fusion pays where the output does not fit in L2 (126 MB): below that the separate statistics
# pass reads the activation from L2 right after the convolution wrote it, and costs little.
_launches = [0]  # kernel launches issued from this module (the trainer reports launches per step)


def launch_count() -> int:
    return _launches[0]


# Input-gradient GEMMs can read the weight [Cout, Cin] as stored (MN-major B operand, ops/gemm.py) instead of a
# transposed copy.  Measured on the ResNet-50 step this LOSES 0.35 ms (19.9 -> 20.3 ms: 17 small transposes saved, but
# every dgrad GEMM of the K = 64..512 layers runs slower with the MN-major operand), so it is off here and on for
# the Wide&Deep tower (0.548 -> 0.543 ms).  PDJ_CONV_KN=1 enables it.
CONV_KN = os.environ.get("PDJ_CONV_KN", "0") == "1"
# backward of a 1x1 convolution WITHOUT a skip gradient on the own kernels (dgrad GEMM + MN-major wgrad) when the input
# is at least this large; smaller layers stay in the library (A/B: profiles/r2/conv1x1_bwd_ab.md)
OWN_BWD = os.environ.get("PDJ_CONV1X1_OWN_BWD", "1") == "1"
OWN_BWD_MIN_BYTES = int(float(os.environ.get("PDJ_CONV1X1_OWN_BWD_MIN_MB", "0")) * (1 << 20))
MIN_OUTPUT_BYTES = int(float(os.environ.get("PDJ_CONV1X1_MIN_MB", "150")) * (1 << 20))


def conv1x1_eligible(x: torch.Tensor, weight: torch.Tensor, stride=1, padding=0, min_bytes: int = 0) -> bool:
    """True when the hand-written kernel can take this convolution."""
    if min_bytes and x.numel() // x.shape[1] * weight.shape[0] * 2 < min_bytes:
        return False
    s = stride if isinstance(stride, int) else (stride[0] if stride[0] == stride[1] else -1)
    p = padding if isinstance(padding, int) else (padding[0] if padding[0] == padding[1] else -1)
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 4
            and weight.dim() == 4 and weight.shape[2] == 1 and weight.shape[3] == 1 and s == 1 and p == 0
            and weight.shape[0] % 64 == 0 and weight.shape[1] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last))


def conv1x1_gemm(x2: torch.Tensor, w2: torch.Tensor, stats: bool = True,
                 out: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """out[M,N] = x2[M,K] @ w2[N,K]^T on the kernel; also the [rows, 2, N] fp32 partial sums when `stats`."""
    lib = _lib.load()
    m, k = x2.shape
    n = w2.shape[0]
    assert x2.stride(1) == 1 and w2.stride(1) == 1 and w2.shape[1] == k
    if out is None:
        out = torch.empty((m, n), dtype=torch.bfloat16, device=x2.device)
    partial = None
    if stats:
        rows = lib.pdj_conv1x1_partial_rows(m, n)
        if rows < 1:
            raise ValueError(f"conv1x1: unsupported shape M={m} N={n}")
        partial = torch.empty((rows, 2, n), dtype=torch.float32, device=x2.device)
    _lib.check(lib.pdj_conv1x1_bf16(x2.data_ptr(), w2.data_ptr(), out.data_ptr(), _lib.ptr(partial), m, n, k,
                                    x2.stride(0), w2.stride(0), out.stride(0),
                                    torch.cuda.current_stream().cuda_stream), "conv1x1_bf16")
    _launches[0] += 1
    return out, partial


def conv1x1_gemm_add(a2: torch.Tensor, b2: torch.Tensor, addend2: Optional[torch.Tensor],
                     out: Optional[torch.Tensor] = None, b_kn: bool = False) -> torch.Tensor:
    """out[M,N] = addend2[M,N] + a2[M,K] @ b2[N,K]^T on the kernel (out may be addend2 itself).  With `b_kn`, b2 is
    [K,N] (a weight [Cout, Cin] used for its input gradient, read as stored) and addend2 may be None."""
    if b_kn and not CONV_KN:
        b2, b_kn = b2.t().contiguous(), False
    m, k = a2.shape
    n = b2.shape[1] if b_kn else b2.shape[0]
    assert a2.stride(1) == 1 and b2.stride(1) == 1 and (b2.shape[0] if b_kn else b2.shape[1]) == k
    assert addend2 is None or (addend2.shape == (m, n) and addend2.stride(1) == 1)
    if out is None:
        out = torch.empty((m, n), dtype=torch.bfloat16, device=a2.device)
    lib = _lib.load()
    if b_kn:
        _lib.check(lib.pdj_conv1x1_kn_bf16(a2.data_ptr(), b2.data_ptr(), _lib.ptr(addend2), out.data_ptr(), m, n, k,
                                           a2.stride(0), b2.stride(0), 0 if addend2 is None else addend2.stride(0),
                                           out.stride(0), torch.cuda.current_stream().cuda_stream), "conv1x1_kn_bf16")
    elif addend2 is None:
        return conv1x1_gemm(a2, b2, stats=False, out=out)[0]
    else:
        _lib.check(lib.pdj_conv1x1_add_bf16(a2.data_ptr(), b2.data_ptr(), addend2.data_ptr(), out.data_ptr(), m, n, k,
                                            a2.stride(0), b2.stride(0), addend2.stride(0), out.stride(0),
                                            torch.cuda.current_stream().cuda_stream), "conv1x1_add_bf16")
    _launches[0] += 1
    return out


def _rows(t: torch.Tensor) -> torch.Tensor:
    """[N,C,H,W] channels-last -> its storage as [N*H*W, C] (a view)."""
    n, c, h, w = t.shape
    v = t.permute(0, 2, 3, 1).reshape(n * h * w, c)
    assert v.data_ptr() == t.data_ptr()
    return v


def _wgrad_1x1(dy: torch.Tensor, x: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """dW[Cout,Cin,1,1] = dY^T[Cout,M] @ X[M,Cin] on the MN-major tcgen05 kernel (ops.gemm.wgrad_mn)."""
    from .gemm import wgrad_mn

    _launches[0] += 2
    sink = sink_of(weight)  # [Cout,Cin,1,1] channels-last: the same dense [Cout, Cin] matrix in memory
    if sink is not None:
        wgrad_mn(_rows(dy), _rows(x), out=sink)
        return adopt(sink)
    return wgrad_mn(_rows(dy), _rows(x)).view(weight.shape)


class _Conv1x1StatsFn(torch.autograd.Function):
    """y = conv1x1(x, weight) with BN partial sums; with `carry_skip` the input is also handed through as a
    third output, so that the gradient of whatever consumes it (the block's skip connection) arrives in
    the same backward call and is folded into the input-gradient GEMM instead of a separate add pass."""

    @staticmethod
    def forward(ctx, x, weight, carry_skip, use_gemm):
        n, c, h, w = x.shape
        o = weight.shape[0]
        if use_gemm:
            y = torch.empty((n, o, h, w), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
            _, partial = conv1x1_gemm(_rows(x), weight.reshape(o, c), stats=True, out=_rows(y))
            ctx.mark_non_differentiable(partial)
        else:  # library forward (small layers); the fused skip-gradient backward below still applies
            y, partial = F.conv2d(x, weight), None
        ctx.save_for_backward(x, weight)
        ctx.carry_skip = carry_skip
        if carry_skip:
            return y, partial, x.view_as(x)
        return y, partial

    @staticmethod
    def backward(ctx, dy, _dpartial, dskip=None):
        x, weight = ctx.saved_tensors
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        o, c = weight.shape[0], weight.shape[1]
        fuse = (ctx.carry_skip and dskip is not None and ctx.needs_input_grad[0] and c % 64 == 0 and o % 8 == 0)
        own = (not fuse and OWN_BWD and c % 64 == 0 and o % 8 == 0 and x.numel() * 2 >= OWN_BWD_MIN_BYTES
               and x.is_contiguous(memory_format=torch.channels_last))
        if fuse:
            # weight gradient as a plain library GEMM dY^T @ X (a wgrad-only cuDNN call picks a far slower kernel)
            dx, dw = None, (_wgrad_1x1(dy, x, weight) if ctx.needs_input_grad[1] else None)
        elif own:
            # no skip gradient to fold in (conv3 of a bottleneck): the same two kernels without the addend
            dw = _wgrad_1x1(dy, x, weight) if ctx.needs_input_grad[1] else None
            dx = None
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x, memory_format=torch.channels_last)
                conv1x1_gemm_add(_rows(dy), weight.reshape(o, c), None, out=_rows(dx), b_kn=True)
        else:
            dx, dw, _ = torch.ops.aten.convolution_backward(
                dy, x, weight, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                [ctx.needs_input_grad[0], ctx.needs_input_grad[1], False])
        if fuse:
            if not dskip.is_contiguous(memory_format=torch.channels_last):
                dskip = dskip.contiguous(memory_format=torch.channels_last)
            # dX[M,Cin] = dskip + dY[M,Cout] @ W[Cout,Cin]: the B operand is the weight as stored (MN-major), no W^T copy
            dx = torch.empty_like(x, memory_format=torch.channels_last)
            conv1x1_gemm_add(_rows(dy), weight.reshape(o, c), _rows(dskip), out=_rows(dx), b_kn=True)
        elif ctx.carry_skip and dskip is not None and dx is not None:
            dx = dx + dskip
        elif ctx.carry_skip and dskip is not None:
            dx = dskip
        return dx, dw, None, None


def conv1x1_stats(x: torch.Tensor, weight: torch.Tensor, carry_skip: bool = False):
    """1x1 / stride 1 convolution; returns (y, partial sums for the following BatchNorm) or, with
    `carry_skip`, (y, partial, x') where x' is x for the skip connection (see _Conv1x1StatsFn).
    `partial` is None where the forward ran in the library (shapes the kernel does not take, or outputs
    below MIN_OUTPUT_BYTES); the fused skip-gradient backward is used whenever the transposed GEMM fits."""
    fwd = conv1x1_eligible(x, weight, min_bytes=MIN_OUTPUT_BYTES)
    bwd = carry_skip and _dgrad_eligible(x, weight)
    if fwd or bwd:
        out = _Conv1x1StatsFn.apply(x, weight, bool(bwd), bool(fwd))
        return (out[0], out[1], x) if (carry_skip and not bwd) else out
    y = F.conv2d(x, weight)
    return (y, None, x) if carry_skip else (y, None)


class _Conv1x1DualFn(torch.autograd.Function):
    """The two 1x1 convolutions that read the input of a down-sampling bottleneck: conv1 (stride 1) and the
    projection shortcut (stride s).  Their input gradients are summed inside one GEMM: the shortcut's
    dgrad comes from the library, conv1's dgrad runs on the tcgen05 kernel with that tensor as addend."""

    @staticmethod
    def forward(ctx, x, w1, w_ds, stride, gemm1, gemm_ds):
        n, c, h, w = x.shape
        o1 = w1.shape[0]
        if gemm1:
            y1 = torch.empty((n, o1, h, w), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
            _, p1 = conv1x1_gemm(_rows(x), w1.reshape(o1, c), stats=True, out=_rows(y1))
            ctx.mark_non_differentiable(p1)
        else:
            y1, p1 = F.conv2d(x, w1), None
        if gemm_ds:
            od = w_ds.shape[0]
            yd = torch.empty((n, od, h, w), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
            _, pd = conv1x1_gemm(_rows(x), w_ds.reshape(od, c), stats=True, out=_rows(yd))
            ctx.mark_non_differentiable(pd)
        else:
            yd, pd = F.conv2d(x, w_ds, stride=stride), None
        ctx.save_for_backward(x, w1, w_ds)
        ctx.stride = stride
        return y1, p1, yd, pd

    @staticmethod
    def backward(ctx, dy1, _dp1, dyd, _dpd):
        x, w1, w_ds = ctx.saved_tensors
        cl = torch.channels_last
        dy1 = dy1 if dy1.is_contiguous(memory_format=cl) else dy1.contiguous(memory_format=cl)
        dyd = dyd if dyd.is_contiguous(memory_format=cl) else dyd.contiguous(memory_format=cl)
        s = ctx.stride
        need_x = ctx.needs_input_grad[0]
        g, dwd, _ = torch.ops.aten.convolution_backward(dyd, x, w_ds, None, [s, s], [0, 0], [1, 1], False, [0, 0], 1,
                                                        [need_x, ctx.needs_input_grad[2], False])
        dw1 = _wgrad_1x1(dy1, x, w1) if ctx.needs_input_grad[1] else None
        dx = None
        if need_x:
            o1, c = w1.shape[0], w1.shape[1]
            g = g if g.is_contiguous(memory_format=cl) else g.contiguous(memory_format=cl)
            conv1x1_gemm_add(_rows(dy1), w1.reshape(o1, c), _rows(g), out=_rows(g), b_kn=True)  # in place: g += dY1 @ W1
            dx = g
        return dx, dw1, dwd, None, None, None


def conv1x1_dual(x: torch.Tensor, w1: torch.Tensor, w_ds: torch.Tensor, stride: int):
    """(conv1(x), partial1, shortcut(x), partial_ds) for a down-sampling bottleneck; None partials where the
    forward ran in the library.  Falls back to two plain convolutions when conv1's dgrad GEMM does not fit."""
    if _dgrad_eligible(x, w1) and w_ds.dim() == 4 and w_ds.shape[2] == 1 and w_ds.shape[3] == 1:
        gemm1 = conv1x1_eligible(x, w1, min_bytes=MIN_OUTPUT_BYTES)
        gemm_ds = stride == 1 and conv1x1_eligible(x, w_ds, min_bytes=MIN_OUTPUT_BYTES)
        return _Conv1x1DualFn.apply(x, w1, w_ds, int(stride), bool(gemm1), bool(gemm_ds))
    y1, p1 = conv1x1_stats(x, w1)
    if stride == 1:
        yd, pd = conv1x1_stats(x, w_ds)
    else:
        yd, pd = F.conv2d(x, w_ds, stride=stride), None
    return y1, p1, yd, pd


def _dgrad_eligible(x: torch.Tensor, weight: torch.Tensor) -> bool:
    """dX = dskip + dY @ W as a GEMM with N = Cin, K = Cout."""
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 4
            and weight.dim() == 4 and weight.shape[2] == 1 and weight.shape[3] == 1
            and weight.shape[1] % 64 == 0 and weight.shape[0] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last))


# ---- ResNet stem: 4x4 / stride-1 convolution over the space-to-depth input -------------------------------
def stem_s2d_eligible(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 4
            and tuple(weight.shape) == (64, 16, 4, 4) and x.shape[1] == 16 and 4 <= x.shape[3] <= 131
            and x.is_contiguous(memory_format=torch.channels_last)
            and weight.is_contiguous(memory_format=torch.channels_last))


class _StemS2dFn(torch.autograd.Function):
    """y = conv2d(x, weight) for the [64,16,4,4] stem over x[N,16,Ho+3,Wo+3] on the tcgen05 pipeline
    (csrc/conv_tcgen05.cu, WIN mode), with the BN partial sums from its epilogue."""

    @staticmethod
    def forward(ctx, x, weight):
        lib = _lib.load()
        n, _, hi, wi = x.shape
        ho, wo = hi - 3, wi - 3
        y = torch.empty((n, 64, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        partial = torch.empty((lib.pdj_stem_s2d_partial_rows(n, ho), 2, 64), dtype=torch.float32, device=x.device)
        _lib.check(lib.pdj_stem_s2d_fwd(x.data_ptr(), weight.data_ptr(), y.data_ptr(), partial.data_ptr(), n, ho, wo,
                                        torch.cuda.current_stream().cuda_stream), "stem_s2d_fwd")
        _launches[0] += 1
        ctx.save_for_backward(x, weight)
        ctx.mark_non_differentiable(partial)
        return y, partial

    @staticmethod
    def backward(ctx, dy, _dpartial):
        x, weight = ctx.saved_tensors
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = dw = None
        if ctx.needs_input_grad[0]:  # not the case in training (x is the image batch)
            dx, _, _ = torch.ops.aten.convolution_backward(dy, x, weight, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1,
                                                           [True, False, False])
        if ctx.needs_input_grad[1]:
            lib = _lib.load()
            n, _, hi, wi = x.shape
            dw = torch.empty_like(weight, memory_format=torch.channels_last)
            ws = torch.empty(lib.pdj_stem_s2d_wgrad_ws_bytes(), dtype=torch.uint8, device=x.device)
            _lib.check(lib.pdj_stem_s2d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), ws.data_ptr(), n, hi - 3, wi - 3,
                                              torch.cuda.current_stream().cuda_stream), "stem_s2d_wgrad")
            _launches[0] += 2  # the tcgen05 kernel and the reduction of its per-CTA partials
        return dx, dw


def stem_s2d_conv(x: torch.Tensor, weight: torch.Tensor):
    """(y, partial) for the space-to-depth stem; library convolution (partial = None) when not eligible."""
    if stem_s2d_eligible(x, weight):
        return _StemS2dFn.apply(x, weight)
    return F.conv2d(x, weight), None


# ---- 3x3 / pad 1 / stride 1 convolution: weight gradient on tcgen05 -------------------------------------------
def conv3x3_wgrad_eligible(x: torch.Tensor, weight: torch.Tensor, stride=1, padding=1) -> bool:
    s = stride if isinstance(stride, int) else (stride[0] if stride[0] == stride[1] else -1)
    p = padding if isinstance(padding, int) else (padding[0] if padding[0] == padding[1] else -1)
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 4
            and weight.dim() == 4 and weight.shape[2] == 3 and weight.shape[3] == 3 and s == 1 and p == 1
            and weight.shape[0] % 64 == 0 and weight.shape[1] % 32 == 0 and x.shape[3] <= 254
            and x.is_contiguous(memory_format=torch.channels_last))


def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor, weight_shape, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dW (channels-last [Cout, Cin, 3, 3] bf16) of y = conv2d(x, w, padding=1) from x and dy, both NHWC bf16
    (csrc/wgrad_tcgen05.cu: one halo tile per pixel block serves all nine taps, accumulators resident in TMEM)."""
    lib = _lib.load()
    n, c, h, w = x.shape
    o = weight_shape[0]
    assert dy.shape == (n, o, h, w) and dy.is_contiguous(memory_format=torch.channels_last)
    if out is not None:  # e.g. a gradient sink: [Cout, Cin, 3, 3] with channels-last strides
        assert out.dtype == torch.bfloat16 and out.shape == (o, c, 3, 3) and out.stride() == (9 * c, 1, 3 * c, c)
        dw = out
    else:
        dw = torch.empty((o, 3, 3, c), dtype=torch.bfloat16, device=x.device).permute(0, 3, 1, 2)  # channels-last strides
    nbytes = lib.pdj_conv3x3_wgrad_ws_bytes(n, h, w, c, o)
    if nbytes == 0:
        raise ValueError(f"conv3x3_wgrad: unsupported shape {tuple(x.shape)} -> {o}")
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    _lib.check(lib.pdj_conv3x3_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), ws.data_ptr(), n, h, w, c, o,
                                     torch.cuda.current_stream().cuda_stream), "conv3x3_wgrad")
    _launches[0] += 2  # the tcgen05 kernel and the reduction of its per-CTA partials
    return dw


def conv_im2col_eligible(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and x.dim() == 4
            and weight.dim() == 4 and weight.shape[2] == weight.shape[3] and weight.shape[2] <= 7
            and weight.shape[0] % 64 == 0 and weight.shape[1] % 64 == 0 and x.shape[1] == weight.shape[1]
            and x.is_contiguous(memory_format=torch.channels_last)
            and weight.is_contiguous(memory_format=torch.channels_last))


def conv_im2col(x: torch.Tensor, weight: torch.Tensor, stride: int = 1, padding: int = 0, stats: bool = False):
    """(y, partial) = conv2d(x, weight, stride, padding) as an implicit GEMM on the tcgen05 pipeline (im2col-mode TMA
    loads for the A operand, csrc/conv_tcgen05.cu); `partial` are the BN partial sums of y when `stats`, else None."""
    lib = _lib.load()
    n, c, h, w = x.shape
    o, k = weight.shape[0], weight.shape[2]
    p_, q_ = (h + 2 * padding - k) // stride + 1, (w + 2 * padding - k) // stride + 1
    y = torch.empty((n, p_, q_, o), dtype=torch.bfloat16, device=x.device).permute(0, 3, 1, 2)
    partial = None
    if stats:
        rows = lib.pdj_conv_im2col_partial_rows(n, h, w, o, k, padding, stride)
        partial = torch.empty((rows, 2, o), dtype=torch.float32, device=x.device)
    _lib.check(lib.pdj_conv_im2col_fwd(x.data_ptr(), weight.data_ptr(), y.data_ptr(), _lib.ptr(partial), n, h, w, c, o, k,
                                       padding, stride, torch.cuda.current_stream().cuda_stream), "conv_im2col_fwd")
    _launches[0] += 1
    return y, partial


def conv3x3_fprop(x: torch.Tensor, weight: torch.Tensor, stride: int = 1, stats: bool = False):
    y, partial = conv_im2col(x, weight, stride, 1, stats)
    return (y, partial) if stats else y


def flipped_transposed(weight: torch.Tensor) -> torch.Tensor:
    """[O, I, k, k] -> the [I, O, k, k] weight whose stride-1 convolution over dY is the input gradient:
    Wd[i, o, r, s] = W[o, i, k-1-r, k-1-s]; channels-last memory [I][k][k][O]."""
    return weight.flip(2, 3).permute(1, 0, 2, 3).contiguous(memory_format=torch.channels_last)


def conv3x3_dgrad(dy: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """dX of y = conv2d(x, weight, padding=1) (stride 1): the same implicit GEMM over dY."""
    return conv_im2col(dy, flipped_transposed(weight), 1, 1, False)[0]


# Which 3x3 layers take the im2col kernels, by input channels.  Measured at batch 256 (profiles/conv3x3_bench.py, and
# end to end in the training step, profiles/r2/conv3x3_ab.md): the 1-CTA 128 x N tile needs ~125 B/clk/SM from L2
# and reaches 0.85 PFLOP/s where the library's 2-CTA tiles reach 1.15, and the statistics pass it saves reads from
# L2 at these sizes, so the step is 0.14 - 0.55 ms FASTER with the library forward.  The kernels stay available
# (tests, PDJ_CONV3X3_FPROP_MIN_C=128 / PDJ_CONV3X3_DGRAD_MIN_C=128) and off by default until they have 2-CTA tiles.
IM2COL_FPROP_MIN_C = int(os.environ.get("PDJ_CONV3X3_FPROP_MIN_C", "100000"))
IM2COL_DGRAD_MIN_C = int(os.environ.get("PDJ_CONV3X3_DGRAD_MIN_C", "100000"))


class _Conv3x3Fn(torch.autograd.Function):
    """(y, partial) = conv2d(x, weight, stride, padding=1) with BN partial sums.  Forward and (stride 1) input gradient
    on the im2col implicit GEMM, the (stride 1) weight gradient on the halo-tile kernel; what a kernel does not
    take runs in the library."""

    @staticmethod
    def forward(ctx, x, weight, stride):
        ctx.save_for_backward(x, weight)
        ctx.stride = stride
        if weight.shape[1] >= IM2COL_FPROP_MIN_C and conv_im2col_eligible(x, weight):
            y, partial = conv_im2col(x, weight, stride, 1, True)
            ctx.mark_non_differentiable(partial)
            return y, partial
        return F.conv2d(x, weight, stride=stride, padding=1), None

    @staticmethod
    def backward(ctx, dy, _dpartial):
        x, weight = ctx.saved_tensors
        s = ctx.stride
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        ours_x = (need_x and s == 1 and weight.shape[0] >= IM2COL_DGRAD_MIN_C and weight.shape[0] % 64 == 0
                  and weight.shape[1] % 64 == 0)
        ours_w = (need_w and s == 1 and conv3x3_wgrad_eligible(x, weight)
                  and min(x.shape[2], x.shape[3]) >= WGRAD3X3_MIN_HW)
        dx = conv3x3_dgrad(dy, weight) if ours_x else None
        dw = None
        if ours_w:
            sink = sink_of(weight)
            dw = conv3x3_wgrad(x, dy, weight.shape, out=sink)
            dw = adopt(dw) if sink is not None else dw
        lib_x, lib_w = need_x and not ours_x, need_w and not ours_w
        if lib_x or lib_w:
            gx, gw, _ = torch.ops.aten.convolution_backward(dy, x, weight, None, [s, s], [1, 1], [1, 1], False, [0, 0], 1,
                                                            [lib_x, lib_w, False])
            dx = gx if lib_x else dx
            dw = gw if lib_w else dw
        return dx, dw, None


WGRAD3X3_MIN_HW = int(os.environ.get("PDJ_WGRAD3X3_MIN_HW", "10"))


def conv3x3(x: torch.Tensor, weight: torch.Tensor, stride: int = 1):
    """(y, partial) for a 3x3 / pad 1 convolution followed by BatchNorm; partial is None where the forward ran in the
    library.  Images smaller than WGRAD3X3_MIN_HW pixels on a side (the 7x7 stage) keep the library weight gradient:
    a 49-pixel image fills 64-row blocks only to 77 % and 64 (o, c) tiles leave no room for an even split (measured
    85 vs 42 us at batch 256)."""
    if (x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and weight.shape[2] == 3
            and x.is_contiguous(memory_format=torch.channels_last)):
        return _Conv3x3Fn.apply(x, weight, int(stride))
    return F.conv2d(x, weight, stride=stride, padding=1), None
