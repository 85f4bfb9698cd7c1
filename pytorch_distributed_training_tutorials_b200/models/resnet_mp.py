"""ResNet-50 split across two GPUs: naive and pipelined.

``ModelParallelResNet50`` uses the reference's split point (03.model_parallel.ipynb:807-833):
``seq1 = conv1, bn1, relu, maxpool, layer1, layer2`` on the first device, ``seq2 = layer3, layer4,
avgpool`` and ``fc`` on the second; 25 557 032 parameters either way (ibid. 857, 888).  Unlike the
reference, stages are not registered twice, so ``state_dict()`` has exactly torchvision's keys.
``PipelineParallelResNet50`` runs the same split on micro-batches of ``split_size`` images.
"""
from __future__ import annotations

import torch
from torch import nn

from ..parallel.model_parallel import Pipeline, hop
from .resnet import ResNet


class _Stage0(nn.Module):
    def __init__(self, net: ResNet) -> None:
        super().__init__()
        self.net = [net]                      # not registered: parameters stay owned by `net`

    def forward(self, x):
        return self.net[0].features_stage0(x)


class _Stage1(nn.Module):
    def __init__(self, net: ResNet) -> None:
        super().__init__()
        self.net = [net]

    def forward(self, x):
        n = self.net[0]
        return n.fc(n.features_stage1(x))


class ModelParallelResNet50(ResNet):
    def __init__(self, num_classes: int = 1000, dev0="cuda:0", dev1="cuda:1") -> None:
        super().__init__([3, 4, 6, 3], num_classes=num_classes)
        self.dev0, self.dev1 = torch.device(dev0), torch.device(dev1)
        for m in (self.conv1, self.bn1, self.maxpool, self.layer1, self.layer2):
            m.to(self.dev0)
        for m in (self.layer3, self.layer4, self.avgpool, self.fc):
            m.to(self.dev1)

    def forward(self, x):
        x = self.features_stage0(x.to(self.dev0))
        x = self.features_stage1(hop(x, self.dev1))
        return self.fc(x)


class PipelineParallelResNet50(ModelParallelResNet50):
    def __init__(self, split_size: int = 20, **kw) -> None:
        super().__init__(**kw)
        self.split_size = split_size
        self._s0, self._s1 = _Stage0(self), _Stage1(self)

    def forward(self, x):
        micro = list(x.to(self.dev0).split(self.split_size, dim=0))
        outs = []
        prev = None
        # fill-drain over two stages: stage 1 consumes micro-batch i-1 while stage 0 produces i
        for mb in micro:
            cur = hop(self._s0(mb), self.dev1)
            if prev is not None:
                outs.append(self._s1(prev))
            prev = cur
        outs.append(self._s1(prev))
        return torch.cat(outs, dim=0)
