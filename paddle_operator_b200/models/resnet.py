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

    def __init__(self, inplanes: int, planes: int, stride: int, fused: bool):
        super().__init__()
        out = planes * self.expansion
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
        y = self.bn1(self.conv1(x))
        y = self.bn2(self.conv2(y))
        if self.downsample is not None:
            identity = self.bn_ds(self.downsample(x))
        return self.bn3(self.conv3(y), identity)


class ResNet(nn.Module):
    def __init__(self, layers=(3, 4, 6, 3), num_classes: int = 1000, fused: bool = True, width: int = 64):
        super().__init__()
        self.fused = fused
        self.conv1 = nn.Conv2d(3, width, 7, stride=2, padding=3, bias=False)
        self.bn1 = _bn(width, True, fused)
        self.maxpool = FusedMaxPool2d(3, 2, 1) if fused else nn.MaxPool2d(3, stride=2, padding=1)
        blocks = []
        inplanes = width
        for i, n in enumerate(layers):
            planes = width * (2 ** i)
            for j in range(n):
                blocks.append(Bottleneck(inplanes, planes, (1 if i == 0 else 2) if j == 0 else 1, fused))
                inplanes = planes * Bottleneck.expansion
        self.blocks = nn.Sequential(*blocks)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(inplanes, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        x = self.bn1(self.conv1(x))
        x = self.maxpool(x)
        x = self.blocks(x)
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)


def resnet50(num_classes: int = 1000, fused: bool = True) -> ResNet:
    return ResNet((3, 4, 6, 3), num_classes, fused)


def resnet_tiny(num_classes: int = 16, fused: bool = True) -> ResNet:
    """Two-stage, narrow variant for smoke tests."""
    return ResNet((1, 1), num_classes, fused, width=16)


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
