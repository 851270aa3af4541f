"""Fused BatchNorm (+residual)(+ReLU) for channels-last bf16 activations.

CUDA path: csrc/bn_act.cu (two kernels forward, two backward).  The pure
PyTorch function `bn_act_reference` is the fp32 oracle used by the tests and
the CPU path of the module.

Reference parity: the reference contains no model code; the workload is the
ResNet of deploy/examples/resnet.yaml:13-19 which Paddle executes inside the
user image.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._sink import adopt, sink_of

_workspaces: dict[int, torch.Tensor] = {}


def _workspace(device: torch.device) -> torch.Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    ws = _workspaces.get(idx)
    if ws is None:
        n = _lib.load().pdj_bn_workspace_bytes()
        ws = torch.zeros(n, dtype=torch.uint8, device=device)
        _workspaces[idx] = ws
    return ws


def bn_act_reference(x, weight, bias, residual=None, relu=True, eps=1e-5):
    """fp32 training-mode BN over (N,H,W) + optional residual + optional ReLU."""
    xf = x.float()
    mean = xf.mean(dim=(0, 2, 3), keepdim=True)
    var = xf.var(dim=(0, 2, 3), unbiased=False, keepdim=True)
    y = (xf - mean) * torch.rsqrt(var + eps) * weight.float().view(1, -1, 1, 1) + bias.float().view(1, -1, 1, 1)
    if residual is not None:
        y = y + residual.float()
    if relu:
        y = torch.relu(y)
    return y


def _is_nhwc(t: torch.Tensor) -> bool:
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last)


class _BnActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual, running_mean, running_var, relu, eps, momentum, w32, b32, partial=None):
        lib = _lib.load()
        assert x.is_cuda and x.dtype == torch.bfloat16
        if not _is_nhwc(x):
            x = x.contiguous(memory_format=torch.channels_last)
        if residual is not None and not _is_nhwc(residual):
            residual = residual.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        m = n * h * w
        y = torch.empty_like(x, memory_format=torch.channels_last)
        stats = torch.empty(4, c, dtype=torch.float32, device=x.device)  # mean, rstd, scale, shift
        # fp32 views of the affine parameters: the trainer's master copies when
        # it provides them, else a cast of the (possibly bf16) parameters.
        gamma = w32 if w32 is not None else (weight if weight.dtype == torch.float32 else weight.float())
        beta = b32 if b32 is not None else (bias if bias.dtype == torch.float32 else bias.float())
        ws = _workspace(x.device)
        st = torch.cuda.current_stream().cuda_stream
        # ReLU derivative for the backward pass as one bit per element (instead of re-reading y)
        gate = torch.empty(m * c // 8, dtype=torch.uint8, device=x.device) if relu else None
        if partial is not None:
            # the producer of x (ops.conv.conv1x1_stats) already reduced the statistics to a few rows
            assert partial.dtype == torch.float32 and partial.dim() == 3 and partial.shape[1:] == (2, c)
            _lib.check(lib.pdj_bn_fwd_pre(x.data_ptr(), _lib.ptr(residual), y.data_ptr(), _lib.ptr(gate), m, c,
                                          gamma.data_ptr(), beta.data_ptr(), eps, momentum, _lib.ptr(running_mean),
                                          _lib.ptr(running_var), stats.data_ptr(), stats[2:].data_ptr(), int(relu),
                                          partial.data_ptr(), partial.shape[0], st), "bn_fwd_pre")
        else:
            _lib.check(lib.pdj_bn_fwd_gate(x.data_ptr(), _lib.ptr(residual), y.data_ptr(), _lib.ptr(gate), m, c,
                                           gamma.data_ptr(), beta.data_ptr(), eps, momentum, _lib.ptr(running_mean),
                                           _lib.ptr(running_var), stats.data_ptr(), stats[2:].data_ptr(), int(relu),
                                           ws.data_ptr(), st), "bn_fwd")
        ctx.save_for_backward(x, gate, gamma, stats)
        ctx.relu = relu
        ctx.has_res = residual is not None
        ctx.wdtype = weight.dtype
        ctx.affine = (weight, bias)  # parameters: backward only looks at their gradient sinks
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _lib.load()
        x, gate, gamma, stats = ctx.saved_tensors
        if not _is_nhwc(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        n, c, h, w = x.shape
        m = n * h * w
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if (ctx.has_res and ctx.relu) else None
        dgb = torch.empty(2, c, dtype=torch.float32, device=x.device)
        dgb16 = torch.empty(2, c, dtype=torch.bfloat16, device=x.device) if ctx.wdtype == torch.bfloat16 else None
        # bf16 gradients of gamma / beta go straight into the trainer's gradient buffer when it provides sinks
        gsink, bsink = sink_of(ctx.affine[0]), sink_of(ctx.affine[1])
        sunk = dgb16 is not None and gsink is not None and bsink is not None
        dg16, db16 = (gsink, bsink) if sunk else ((dgb16[0], dgb16[1]) if dgb16 is not None else (None, None))
        coef = torch.empty(3, c, dtype=torch.float32, device=x.device)
        ws = _workspace(x.device)
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.pdj_bn_bwd_gate2(dy.data_ptr(), None, _lib.ptr(gate), x.data_ptr(), dx.data_ptr(), _lib.ptr(dres),
                                        m, c, gamma.data_ptr(), stats.data_ptr(), dgb.data_ptr(), _lib.ptr(dg16),
                                        _lib.ptr(db16), coef.data_ptr(), int(ctx.relu), ws.data_ptr(), st), "bn_bwd")
        if ctx.has_res and not ctx.relu:
            dres = dy  # no activation: the skip branch sees dy unchanged
        if sunk:
            dgamma, dbeta = adopt(gsink), adopt(bsink)
        elif dgb16 is not None:
            dgamma, dbeta = dgb16[0], dgb16[1]
        else:
            dgamma, dbeta = dgb[0].to(ctx.wdtype), dgb[1].to(ctx.wdtype)
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None, None, None


def bn_act(x, weight, bias, residual: Optional[torch.Tensor] = None, running_mean=None, running_var=None,
           relu: bool = True, eps: float = 1e-5, momentum: float = 0.1, weight_f32=None, bias_f32=None,
           partial: Optional[torch.Tensor] = None):
    """y = relu?(BN_train(x) + residual?).  CUDA bf16 -> hand-written kernels.  `partial` ([rows, 2, C]
    fp32 sums / sums of squares from ops.conv.conv1x1_stats) replaces the statistics pass over x."""
    if x.is_cuda:
        return _BnActFn.apply(x, weight, bias, residual, running_mean, running_var, relu, eps, momentum,
                              weight_f32, bias_f32, partial)
    # CPU path (tests, plumbing): plain PyTorch.
    y = torch.nn.functional.batch_norm(x.float(), running_mean, running_var, weight.float(), bias.float(), True,
                                       momentum, eps)
    if residual is not None:
        y = y + residual.float()
    return (torch.relu(y) if relu else y).to(x.dtype)


class FusedBnAct(torch.nn.Module):
    """BatchNorm2d whose training forward is fused with add and ReLU."""

    def __init__(self, channels: int, relu: bool = True, eps: float = 1e-5, momentum: float = 0.1,
                 zero_init: bool = False):
        super().__init__()
        self.relu, self.eps, self.momentum = relu, eps, momentum
        # Statistics are fp32.  The affine parameters follow the model dtype;
        # when a trainer keeps fp32 masters it points weight_f32/bias_f32 at
        # them so the kernels never see rounded gamma/beta.
        self.weight_f32 = None
        self.bias_f32 = None
        self.weight = torch.nn.Parameter(torch.zeros(channels) if zero_init else torch.ones(channels))
        self.bias = torch.nn.Parameter(torch.zeros(channels))
        self.register_buffer("running_mean", torch.zeros(channels))
        self.register_buffer("running_var", torch.ones(channels))

    def forward(self, x, residual=None, partial=None):
        if self.training:
            return bn_act(x, self.weight, self.bias, residual, self.running_mean, self.running_var, self.relu,
                          self.eps, self.momentum, self.weight_f32, self.bias_f32, partial)
        y = torch.nn.functional.batch_norm(x.float(), self.running_mean, self.running_var, self.weight.float(),
                                           self.bias.float(), False, self.momentum, self.eps)
        if residual is not None:
            y = y + residual.float()
        return (torch.relu(y) if self.relu else y).to(x.dtype)
