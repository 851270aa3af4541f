"""ResNet-50 (v1.5) for the Collective-mode workload.

`fused=True` (the product): cuDNN convolutions + the hand-written fused
BN/add/ReLU kernels of csrc/bn_act.cu.  `fused=False` builds the same network
out of stock torch.nn modules with identical parameter names, which is both the
DDP/NCCL baseline arm of bench.py and the numerical oracle in the tests.

Reference parity: the reference only launches this workload
(deploy/examples/resnet.yaml:13-19: `python -m paddle.distributed.launch
train_fleet.py` with one GPU per worker); the network itself lives in the user
image there.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops.bn import FusedBnAct
from ..ops.conv import conv1x1_dual, conv1x1_stats, conv3x3, stem_s2d_conv
from ..ops.pool import FusedMaxPool2d


class _TorchBnAct(nn.Module):
    """Stock BatchNorm2d (+add)(+ReLU), parameter-compatible with FusedBnAct."""

    def __init__(self, channels: int, relu: bool = True, zero_init: bool = False):
        super().__init__()
        self.relu = relu
        self.bn = nn.BatchNorm2d(channels)
        if zero_init:
            nn.init.zeros_(self.bn.weight)

    # expose the same names as FusedBnAct so state_dicts are interchangeable
    @property
    def weight(self):
        return self.bn.weight

    @property
    def bias(self):
        return self.bn.bias

    def forward(self, x, residual=None):
        y = self.bn(x)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if self.relu else y

    def _load_from_state_dict(self, state_dict, prefix, *a, **k):
        for n in ("weight", "bias", "running_mean", "running_var"):
            if prefix + n in state_dict:
                state_dict[prefix + "bn." + n] = state_dict.pop(prefix + n)
        super()._load_from_state_dict(state_dict, prefix, *a, **k)


def _bn(channels, relu, fused, zero_init=False):
    return FusedBnAct(channels, relu=relu, zero_init=zero_init) if fused else _TorchBnAct(channels, relu, zero_init)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int, fused: bool, gemm_1x1: bool = False):
        super().__init__()
        out = planes * self.expansion
        # stride-1 1x1 convolutions on the tcgen05 GEMM with BN statistics in its epilogue (ops/conv.py)
        self.gemm_1x1 = fused and gemm_1x1
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = _bn(planes, True, fused)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = _bn(planes, True, fused)
        self.conv3 = nn.Conv2d(planes, out, 1, bias=False)
        self.bn3 = _bn(out, True, fused, zero_init=True)  # relu(bn3(.) + identity)
        self.downsample = None
        if stride != 1 or inplanes != out:
            self.downsample = nn.Conv2d(inplanes, out, 1, stride=stride, bias=False)
            self.bn_ds = _bn(out, False, fused)

    def forward(self, x):
        identity = x
        if self.gemm_1x1:
            if self.downsample is None:
                # the skip connection's gradient comes back through conv1's backward (one fused GEMM + add)
                y, p, identity = conv1x1_stats(x, self.conv1.weight, carry_skip=True)
            else:
                # conv1 and the projection shortcut read the same input: their input gradients are summed in
                # conv1's dgrad GEMM
                y, p, d, pd = conv1x1_dual(x, self.conv1.weight, self.downsample.weight, self.downsample.stride[0])
                identity = self.bn_ds(d, partial=pd)
            y = self.bn1(y, partial=p)
            # 3x3 on the tcgen05 kernels: im2col implicit GEMM forward (BN statistics fused) and input gradient,
            # halo-tile weight gradient (csrc/conv_tcgen05.cu, csrc/wgrad_tcgen05.cu)
            y, p = conv3x3(y, self.conv2.weight, self.conv2.stride[0])
            y = self.bn2(y, partial=p)
            y, p = conv1x1_stats(y, self.conv3.weight)
            return self.bn3(y, identity, partial=p)
        y = self.bn1(self.conv1(x))
        y = self.bn2(self.conv2(y))
        if self.downsample is not None:
            identity = self.bn_ds(self.downsample(x))
        return self.bn3(self.conv3(y), identity)


def stem_weight_to_s2d(w7: torch.Tensor) -> torch.Tensor:
    """[O,3,7,7] stem weights -> the [O,16,4,4] weights of the equivalent stride-1 convolution over the
    space-to-depth input (ops.optim.space_to_depth_reference):
        W[o, (dy*2+dx)*3 + c, jy, jx] = w7[o, c, 2*jy + dy - 1, 2*jx + dx - 1]   (zero where out of range)."""
    o = w7.shape[0]
    w8 = w7.new_zeros((o, 3, 8, 8))
    w8[:, :, 1:, 1:] = w7                                   # tap k of the 7x7 kernel sits at k + 1
    w = w8.reshape(o, 3, 4, 2, 4, 2)                          # o c jy dy jx dx
    w = w.permute(0, 3, 5, 1, 2, 4).reshape(o, 12, 4, 4)      # o (dy dx c) jy jx
    out = w7.new_zeros((o, 16, 4, 4))
    out[:, :12] = w
    return out


def stem_s2d_mask(out_channels: int) -> torch.Tensor:
    """1 where an s2d stem weight corresponds to a real 7x7 tap, 0 for the structural zeros."""
    return (stem_weight_to_s2d(torch.ones(out_channels, 3, 7, 7)) != 0).float()


class ResNet(nn.Module):
    def __init__(self, layers=(3, 4, 6, 3), num_classes: int = 1000, fused: bool = True, width: int = 64,
                 s2d_stem: bool = False, gemm_1x1: bool = False):
        super().__init__()
        self.fused = fused
        self.gemm_1x1 = fused and gemm_1x1
        self.s2d_stem = s2d_stem
        if s2d_stem:
            # The 7x7 / stride 2 / pad 3 stem as a 4x4 / stride 1 convolution over the space-to-depth
            # input (16 channels, border baked into the input).  Same function, tensor-core friendly.
            self.conv1 = nn.Conv2d(16, width, 4, stride=1, padding=0, bias=False)
            self.register_buffer("stem_mask", stem_s2d_mask(width), persistent=False)
        else:
            self.conv1 = nn.Conv2d(3, width, 7, stride=2, padding=3, bias=False)
        self.bn1 = _bn(width, True, fused)
        self.maxpool = FusedMaxPool2d(3, 2, 1) if fused else nn.MaxPool2d(3, stride=2, padding=1)
        blocks = []
        inplanes = width
        for i, n in enumerate(layers):
            planes = width * (2 ** i)
            for j in range(n):
                blocks.append(Bottleneck(inplanes, planes, (1 if i == 0 else 2) if j == 0 else 1, fused, gemm_1x1))
                inplanes = planes * Bottleneck.expansion
        self.blocks = nn.Sequential(*blocks)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(inplanes, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        if s2d_stem:
            w7 = torch.empty(width, 3, 7, 7)
            nn.init.kaiming_normal_(w7, mode="fan_out", nonlinearity="relu")
            with torch.no_grad():
                self.conv1.weight.copy_(stem_weight_to_s2d(w7))
            # gradients of the structurally-zero taps are dropped so the stem stays a 7x7 convolution
            self.conv1.weight.register_hook(lambda g: g * self.stem_mask.to(g.dtype))

    def forward(self, x):
        if self.s2d_stem and x.shape[1] == 3:
            from ..ops.optim import space_to_depth_reference

            x = space_to_depth_reference(x).contiguous(memory_format=torch.channels_last)
        if self.s2d_stem and self.gemm_1x1:
            # stem on the tcgen05 pipeline (TMA windows over the space-to-depth input), BN statistics fused
            x, p = stem_s2d_conv(x, self.conv1.weight)
            x = self.bn1(x, partial=p)
        else:
            x = self.bn1(self.conv1(x))
        x = self.maxpool(x)
        x = self.blocks(x)
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)


def resnet50(num_classes: int = 1000, fused: bool = True, s2d_stem: bool = False, gemm_1x1: bool = False) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes, fused, s2d_stem=s2d_stem, gemm_1x1=gemm_1x1)


def resnet_tiny(num_classes: int = 16, fused: bool = True, s2d_stem: bool = False, gemm_1x1: bool = False) -> ResNet:
    """Two-stage, narrow variant for smoke tests (width 16: conv3 of each block is wide enough for the GEMM path)."""
    return ResNet((1, 1), num_classes, fused, width=16, s2d_stem=s2d_stem, gemm_1x1=gemm_1x1)


def to_training_dtype(model: nn.Module, dtype=torch.bfloat16, bn_fp32: bool = False) -> nn.Module:
    """Casts parameters to `dtype` in channels-last layout but keeps the BN
    running statistics in fp32.  `bn_fp32` also keeps stock BatchNorm2d affine
    parameters in fp32, which is what cuDNN's mixed-precision NHWC batch norm
    (the fastest stock path, used by the baseline arm) requires."""
    model = model.to(memory_format=torch.channels_last)
    keep = set()
    if bn_fp32:
        for m in model.modules():
            if isinstance(m, nn.BatchNorm2d):
                keep.update(id(p) for p in m.parameters())
    for p in model.parameters():
        if id(p) not in keep:
            p.data = p.data.to(dtype)
    for name, b in model.named_buffers():
        if b.is_floating_point():
            b.data = b.data.float()
    return model
