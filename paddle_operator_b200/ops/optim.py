"""Python faces of csrc/optim_misc.cu: flat fused SGD, uint8 image
pre-processing, softmax cross-entropy (forward + backward in one kernel).

Each has a PyTorch fp32 reference used by the tests and by the CPU path.
Reference parity: no counterpart in the reference (it has no training code);
see deploy/examples/resnet.yaml:13-19 for the workload these serve.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def sgd_momentum_(master: torch.Tensor, momentum_buf: torch.Tensor, grad_bf16: torch.Tensor,
                  param_bf16: torch.Tensor, hyper: torch.Tensor, wd_split: int) -> None:
    """In place over flat buffers: g = grad*hyper[3] + wd*w (first wd_split
    elements only); v = mu*v + g; w -= lr*v; param = bf16(w).
    hyper = [lr, momentum, weight_decay, grad_scale] lives on the device."""
    n = master.numel()
    if master.is_cuda:
        lib = _lib.load()
        _lib.check(lib.pdj_sgd_momentum(master.data_ptr(), momentum_buf.data_ptr(), grad_bf16.data_ptr(),
                                        param_bf16.data_ptr(), n, wd_split, hyper.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream), "sgd_momentum")
        return
    lr, mu, wd, gs = (float(v) for v in hyper.tolist())
    g = grad_bf16.float() * gs
    g[:wd_split] += wd * master[:wd_split]
    momentum_buf.mul_(mu).add_(g)
    master.add_(momentum_buf, alpha=-lr)
    param_bf16.copy_(master.to(torch.bfloat16))


IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def preprocess_u8_nhwc(images_u8: torch.Tensor, mean=IMAGENET_MEAN, std=IMAGENET_STD, out=None) -> torch.Tensor:
    """uint8 [N,H,W,3] -> bf16 logical [N,3,H,W] in channels-last memory,
    (x/255 - mean)/std.  `out`, if given, is a contiguous bf16 [N,H,W,3] buffer."""
    n, h, w, c = images_u8.shape
    assert c == 3 and images_u8.dtype == torch.uint8 and images_u8.is_contiguous()
    if images_u8.is_cuda:
        if out is None:
            out = torch.empty((n, h, w, 3), dtype=torch.bfloat16, device=images_u8.device)
        assert out.is_contiguous() and out.dtype == torch.bfloat16 and out.numel() == images_u8.numel()
        m = (ctypes.c_float * 3)(*mean)
        s = (ctypes.c_float * 3)(*std)
        _lib.check(_lib.load().pdj_preprocess_u8(images_u8.data_ptr(), out.data_ptr(), n * h * w, m, s,
                                                 torch.cuda.current_stream().cuda_stream), "preprocess_u8")
        return out.permute(0, 3, 1, 2)
    x = images_u8.float() / 255.0
    x = ((x - torch.tensor(mean)) / torch.tensor(std)).to(torch.bfloat16)
    if out is not None:
        out.copy_(x)
        x = out
    return x.permute(0, 3, 1, 2)


class _SoftmaxXentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        rows, ncls = logits.shape
        logits = logits.contiguous()
        dlogits = torch.empty_like(logits)
        loss_sum = torch.zeros(1, dtype=torch.float32, device=logits.device)
        _lib.check(_lib.load().pdj_softmax_xent(logits.data_ptr(), labels.data_ptr(), dlogits.data_ptr(),
                                                loss_sum.data_ptr(), rows, ncls, 1.0 / rows,
                                                torch.cuda.current_stream().cuda_stream), "softmax_xent")
        ctx.save_for_backward(dlogits)
        return loss_sum[0] / rows

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        # g is 1.0 in every training loop here; multiply only when it is not.
        return dlogits * g.to(dlogits.dtype), None


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Mean cross-entropy over rows; bf16 CUDA logits use the fused kernel."""
    if logits.is_cuda and logits.dtype == torch.bfloat16:
        return _SoftmaxXentFn.apply(logits, labels)
    return torch.nn.functional.cross_entropy(logits.float(), labels)


# ---------------------------------------------------------------------------
# Space-to-depth stem (see models/resnet.py): input layout helpers.
S2D_CHANNELS = 16


def s2d_shape(n: int, h: int, w: int) -> tuple:
    """NHWC shape of the stem input for an n x h x w x 3 image batch."""
    return (n, h // 2 + 3, w // 2 + 3, S2D_CHANNELS)


def space_to_depth_reference(x_nchw: torch.Tensor) -> torch.Tensor:
    """Plain-torch definition: [N,3,H,W] -> logical [N,16,H/2+3,W/2+3]; channel (dy*2+dx)*3+c of block
    (by,bx) is pixel (2by+dy, 2bx+dx) channel c; blocks are offset by 2 (zero border 2 before, 1 after)."""
    n, c, h, w = x_nchw.shape
    assert c == 3 and h % 2 == 0 and w % 2 == 0
    x = x_nchw.reshape(n, 3, h // 2, 2, w // 2, 2)            # n c by dy bx dx
    x = x.permute(0, 3, 5, 1, 2, 4).reshape(n, 12, h // 2, w // 2)  # (dy dx c) by bx
    out = x.new_zeros((n, S2D_CHANNELS, h // 2 + 3, w // 2 + 3))
    out[:, :12, 2:2 + h // 2, 2:2 + w // 2] = x
    return out


def preprocess_u8_s2d(images_u8: torch.Tensor, mean=IMAGENET_MEAN, std=IMAGENET_STD, out=None) -> torch.Tensor:
    """uint8 [N,H,W,3] -> normalised bf16 stem input, logical [N,16,H/2+3,W/2+3] in channels-last memory.
    `out`, if given, is a contiguous bf16 [N,H/2+3,W/2+3,16] buffer whose border is already zero."""
    n, h, w, c = images_u8.shape
    assert c == 3 and images_u8.dtype == torch.uint8 and images_u8.is_contiguous()
    if images_u8.is_cuda:
        if out is None:
            out = torch.zeros(s2d_shape(n, h, w), dtype=torch.bfloat16, device=images_u8.device)
        assert out.is_contiguous() and tuple(out.shape) == s2d_shape(n, h, w)
        m = (ctypes.c_float * 3)(*mean)
        s = (ctypes.c_float * 3)(*std)
        _lib.check(_lib.load().pdj_preprocess_u8_s2d(images_u8.data_ptr(), out.data_ptr(), n, h, w, m, s,
                                                     torch.cuda.current_stream().cuda_stream), "preprocess_u8_s2d")
        return out.permute(0, 3, 1, 2)
    x = (images_u8.float() / 255.0 - torch.tensor(mean)) / torch.tensor(std)
    y = space_to_depth_reference(x.permute(0, 3, 1, 2)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    if out is not None:
        out.copy_(y.permute(0, 2, 3, 1))
        return out.permute(0, 3, 1, 2)
    return y
