"""ResNet-50 on this library's sm_100a kernels (NHWC bf16 activations, fp32 master weights).

Same topology, parameter names and parameter count (25 557 032) as torchvision's
``ResNet(Bottleneck, [3, 4, 6, 3])`` that the reference subclasses at
/root/reference/03.model_parallel.ipynb:807-833, so ``state_dict`` files are interchangeable.
Differences are internal: every conv is the tcgen05 implicit GEMM, BN + ReLU (+ residual add) is
one fused kernel pair, and activations never leave NHWC bf16 between the stem and the classifier.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch import nn

from ..ops import functional as Fn
from ..ops import nn as bnn


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: bool = False) -> None:
        super().__init__()
        self.conv1 = bnn.Conv2d(inplanes, planes, 1)
        self.bn1 = bnn.BatchNorm2d(planes, relu=True)
        self.conv2 = bnn.Conv2d(planes, planes, 3, stride=stride)        # stride on the 3x3 (v1.5)
        self.bn2 = bnn.BatchNorm2d(planes, relu=True)
        self.conv3 = bnn.Conv2d(planes, planes * 4, 1)
        self.bn3 = bnn.BatchNorm2d(planes * 4, relu=True)               # relu applied after the add
        self.downsample: Optional[nn.Sequential] = None
        if downsample:
            self.downsample = nn.Sequential(
                bnn.Conv2d(inplanes, planes * 4, 1, stride=stride),
                bnn.BatchNorm2d(planes * 4, relu=False),
            )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        t = self.training            # in training the convs emit the BN statistics from their epilogue
        if t and x.is_cuda and torch.is_grad_enabled():
            # the two gradients of x (first conv + identity / projection) are added inside the dgrad epilogue
            if self.downsample is None:
                c1, s1, identity = Fn.conv2d_skip(x, self.conv1.weight, 1, True)
            else:
                ds = self.downsample[0]
                c1, s1, d, sd = Fn.conv2d_dual(x, self.conv1.weight, ds.weight, 1, ds.stride)
                identity = self.downsample[1](d, sums=sd)
            out = self.bn1(c1, sums=s1)
            o = self.conv2(out, with_stats=True)
            out = self.bn2(o[0], sums=o[1])
            o = self.conv3(out, with_stats=True)
            return self.bn3(o[0], residual=identity, sums=o[1])
        if self.downsample is None:
            identity = x
        else:
            d = self.downsample[0](x, with_stats=t)
            identity = self.downsample[1](d[0], sums=d[1]) if t else self.downsample[1](d)
        o = self.conv1(x, with_stats=t)
        out = self.bn1(o[0], sums=o[1]) if t else self.bn1(o)
        o = self.conv2(out, with_stats=t)
        out = self.bn2(o[0], sums=o[1]) if t else self.bn2(o)
        o = self.conv3(out, with_stats=t)
        return self.bn3(o[0], residual=identity, sums=o[1]) if t else self.bn3(o, residual=identity)


class ResNet(nn.Module):
    def __init__(self, layers: List[int], num_classes: int = 1000, zero_init_residual: bool = False) -> None:
        super().__init__()
        self.inplanes = 64
        self.conv1 = bnn.StemConv(64)
        self.bn1 = bnn.BatchNorm2d(64, relu=True)
        self.maxpool = bnn.MaxPool3x3s2()
        self.layer1 = self._make_layer(64, layers[0], 1)
        self.layer2 = self._make_layer(128, layers[1], 2)
        self.layer3 = self._make_layer(256, layers[2], 2)
        self.layer4 = self._make_layer(512, layers[3], 2)
        self.avgpool = bnn.GlobalAvgPool()
        self.fc = bnn.Linear(512 * 4, num_classes, out_fp32=True)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.zeros_(m.bn3.weight)

    def _make_layer(self, planes: int, blocks: int, stride: int) -> nn.Sequential:
        need_ds = stride != 1 or self.inplanes != planes * 4
        mods = [Bottleneck(self.inplanes, planes, stride, need_ds)]
        self.inplanes = planes * 4
        mods += [Bottleneck(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*mods)

    SEGMENTS = ("stem", "layer1", "layer2", "layer3", "layer4", "pool")

    def run_segment(self, name: str, x: torch.Tensor) -> torch.Tensor:
        if name == "stem":
            if self.training:
                c, s = self.conv1(x, with_stats=True)
                return self.maxpool(self.bn1(c, sums=s))
            return self.maxpool(self.bn1(self.conv1(x)))
        if name == "pool":
            return self.avgpool(x)
        return getattr(self, name)(x)

    def run_segments(self, x: torch.Tensor, first: str, last: str) -> torch.Tensor:
        """Run the consecutive segments ``first .. last`` (inclusive) of SEGMENTS."""
        i, j = self.SEGMENTS.index(first), self.SEGMENTS.index(last)
        for name in self.SEGMENTS[i:j + 1]:
            x = self.run_segment(name, x)
        return x

    # ---- block-level units: the stem, every bottleneck block, the global pool.  Pipeline cuts may fall between
    # any two of them (``"layer2.1"`` = after the second block of layer2); a whole-layer name is its last block.
    def unit_names(self) -> List[str]:
        names = ["stem"]
        for layer in self.SEGMENTS[1:5]:
            names += [f"{layer}.{b}" for b in range(len(getattr(self, layer)))]
        return names + ["pool"]

    def unit_index(self, name: str, last: bool = True) -> int:
        names = self.unit_names()
        if name in names:
            return names.index(name)
        if name in self.SEGMENTS[1:5]:
            return names.index(f"{name}.{len(getattr(self, name)) - 1 if last else 0}")
        raise ValueError(f"unknown unit {name!r}; choose from {names} or a whole layer name")

    def unit_modules(self, name: str) -> List[nn.Module]:
        if name == "stem":
            return [self.conv1, self.bn1, self.maxpool]
        if name == "pool":
            return [self.avgpool]
        layer, b = name.split(".")
        return [getattr(self, layer)[int(b)]]

    def run_unit(self, name: str, x: torch.Tensor) -> torch.Tensor:
        if "." in name:
            layer, b = name.split(".")
            return getattr(self, layer)[int(b)](x)
        return self.run_segment(name, x)

    def run_units(self, x: torch.Tensor, first: int, last: int) -> torch.Tensor:
        """Run units ``first .. last`` (inclusive indices into ``unit_names()``)."""
        for name in self.unit_names()[first:last + 1]:
            x = self.run_unit(name, x)
        return x

    def features_stage0(self, x: torch.Tensor) -> torch.Tensor:
        """conv1 .. layer2 (the reference's ``seq1``, 03.model_parallel.ipynb:812-819)."""
        return self.run_segments(x, "stem", "layer2")

    def features_stage1(self, x: torch.Tensor) -> torch.Tensor:
        """layer3 .. avgpool (the reference's ``seq2``, 03.model_parallel.ipynb:821-825)."""
        return self.run_segments(x, "layer3", "pool")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: ``[N, 3, H, W]`` fp32 (NCHW, like torchvision) -> logits ``[N, num_classes]`` fp32."""
        return self.fc(self.features_stage1(self.features_stage0(x)))


def resnet50(num_classes: int = 1000, **kw) -> ResNet:
    return ResNet([3, 4, 6, 3], num_classes=num_classes, **kw)
