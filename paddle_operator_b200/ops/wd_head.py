"""Fused Wide&Deep head: output layer + wide part + binary cross-entropy, forward and backward.

    loss = head_bce_loss(h, w_out, b_out, wide_sum, dense, w_d, b_d, labels)
         = mean_b BCEWithLogits(h[b] . w_out + b_out + wide_sum[b] + dense[b] . w_d + b_d, labels[b])

CUDA bf16: one kernel computes every sample's logit, loss and dlogit (csrc/optim_misc.cu::wd_head_kernel,
one warp per sample); backward forms the gradient entering the top ReLU layer with the ReLU mask already
applied (wd_head_dz_kernel) - `h` must therefore be the OUTPUT of a ReLU (ops.mlp.FusedMLP with
top_grad_masked=True skips its own mask pass).  All four parameter gradients come from one weighted column-sum
kernel (colsum_w_kernel).
Anything else (CPU, fp32) runs the same math in plain torch.

Reference parity: none - the reference has no model code; the loss of the Wide&Deep workload of
deploy/examples/wide_and_deep.yaml.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import _lib
from ._sink import adopt, sink_of
from .mlp import colsum_w


def head_bce_reference(h, w_out, b_out, wide_sum, dense, w_d, b_d, labels):
    z = F.linear(h, w_out, b_out).squeeze(1).float() + wide_sum.float() \
        + F.linear(dense.to(w_d.dtype), w_d, b_d).squeeze(1).float()
    return F.binary_cross_entropy_with_logits(z, labels.float())


class _HeadBceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, w_out, b_out, wide_sum, dense, w_d, b_d, labels):
        lib = _lib.load()
        bsz, n = h.shape
        dd = dense.shape[1]
        assert h.stride(1) == 1 and w_out.is_contiguous() and w_d.is_contiguous()
        dense32 = dense if dense.dtype == torch.float32 else dense.float()
        dense32 = dense32.contiguous()
        wide32 = wide_sum if wide_sum.dtype == torch.float32 else wide_sum.float()
        lab32 = labels if labels.dtype == torch.float32 else labels.float()
        loss_b = torch.empty(bsz, dtype=torch.float32, device=h.device)
        dlogit = torch.empty(bsz, dtype=torch.float32, device=h.device)
        _lib.check(lib.pdj_wd_head(h.data_ptr(), h.stride(0), w_out.data_ptr(), _lib.ptr(b_out), wide32.contiguous().data_ptr(),
                                   dense32.data_ptr(), w_d.data_ptr(), _lib.ptr(b_d), lab32.contiguous().data_ptr(),
                                   loss_b.data_ptr(), dlogit.data_ptr(), bsz, n, dd, 1.0 / bsz,
                                   torch.cuda.current_stream().cuda_stream), "wd_head")
        ctx.save_for_backward(h, w_out, dense32, dlogit)
        ctx.has = (b_out is not None, b_d is not None)
        ctx.params = (b_out, w_d, b_d)  # parameters: backward only looks at their gradient sinks
        ctx.dtypes = (w_out.dtype, w_d.dtype, wide_sum.dtype)
        return loss_b.mean()

    @staticmethod
    def backward(ctx, g):
        lib = _lib.load()
        h, w_out, dense32, dlogit = ctx.saved_tensors
        bsz, n = h.shape
        dl = dlogit * g  # upstream scalar (1 in training)
        dh = torch.empty_like(h)
        _lib.check(lib.pdj_wd_head_dz(dl.data_ptr(), w_out.data_ptr(), h.data_ptr(), h.stride(0), dh.data_ptr(), dh.stride(0),
                                      bsz, n, torch.cuda.current_stream().cuda_stream), "wd_head_dz")
        # every parameter gradient of the head is a column sum weighted by dlogit: one deterministic kernel, written
        # straight into the trainer's gradient buffer where it provides sinks (ops/_sink.py)
        b_out, w_d, b_d = ctx.params
        dd = dense32.shape[1]
        outs, dests = [], []
        for prm, shape, col, dt in ((w_out, (1, n), 0, ctx.dtypes[0]), (w_d, (1, dd), n, ctx.dtypes[1]),
                                    (b_out, (1,), n + dd, ctx.dtypes[0]), (b_d, (1,), n + dd, ctx.dtypes[1])):
            if prm is None:
                outs.append(None)
                continue
            sink = sink_of(prm, dt)
            t = sink if (sink is not None and sink.is_contiguous()) else torch.empty(shape, dtype=dt, device=h.device)
            dests.append((t, col))
            outs.append(adopt(t) if t is sink else t)
        colsum_w(h, dests, weights=dl, extra=dense32, ones=True)
        dw_out, dw_d, db_out, db_d = outs
        return dh, dw_out, db_out, dl.to(ctx.dtypes[2]), None, dw_d, db_d, None


def head_bce_eligible(h: torch.Tensor, w_out: torch.Tensor, dense: torch.Tensor) -> bool:
    return (h.is_cuda and h.dtype == torch.bfloat16 and w_out.dtype == torch.bfloat16 and h.dim() == 2
            and h.shape[1] % 8 == 0 and h.stride(1) == 1 and h.stride(0) % 8 == 0 and dense.shape[1] <= 32
            and h.data_ptr() % 16 == 0)


def head_bce_loss(h, w_out, b_out, wide_sum, dense, w_d, b_d, labels):
    """Mean BCE-with-logits of the Wide&Deep logit; `h` must be a ReLU output (its positive entries are the
    mask of the gradient returned for it)."""
    if head_bce_eligible(h, w_out, dense) and w_d.dtype == torch.bfloat16:
        return _HeadBceFn.apply(h, w_out, b_out, wide_sum, dense, w_d, b_d, labels)
    return head_bce_reference(h, w_out, b_out, wide_sum, dense, w_d, b_d, labels)
