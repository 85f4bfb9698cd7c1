"""The tutorial's toy models.

* ``ToyLinear``      ``Linear(20, 1)`` regression head of the DDP scripts (ddp_gpus.py:81)
* ``SampleModel``    ``Linear(32, 2)`` that prints its per-replica input shape (01.data_parallel.ipynb:168-175)
* ``ToyModel``       two-device model: ``Linear(10000,10)`` on dev0 -> ReLU -> ``Linear(10,5)`` on dev1
                     (03.model_parallel.ipynb:440-450)
* ``MLP``            a wider regression MLP on the tcgen05 Linear for the DP-vs-DDP throughput config
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import nn as bnn
from ..parallel.model_parallel import hop


class ToyLinear(bnn.Linear):
    """On a GPU this is the fp32 SIMT kernel of csrc/skinny.cu (20 features: no TMA row pitch), on the
    CPU plain ``F.linear``; parameters, initialisation and ``state_dict`` keys are nn.Linear's."""

    def __init__(self) -> None:
        super().__init__(20, 1)


class SampleModel(nn.Module):
    def __init__(self, input_size: int = 32, output_size: int = 2, verbose: bool = True) -> None:
        super().__init__()
        self.fc = bnn.Linear(input_size, output_size)
        self.verbose = verbose

    def forward(self, input):
        output = self.fc(input)
        if self.verbose:
            print("\tIn Model: input size", input.size(), "output size", output.size())
        return output


class ToyModel(nn.Module):
    def __init__(self, dev0="cuda:0", dev1="cuda:1", in_features: int = 10000) -> None:
        super().__init__()
        self.dev0, self.dev1 = torch.device(dev0), torch.device(dev1)
        self.net1 = bnn.Linear(in_features, 10).to(self.dev0)
        self.relu = nn.ReLU()
        self.net2 = bnn.Linear(10, 5).to(self.dev1)

    def forward(self, x):
        x = self.relu(self.net1(x.to(self.dev0)))
        return self.net2(hop(x, self.dev1))


class MLP(nn.Module):
    """Regression MLP ``in -> hidden x depth -> out`` on the tcgen05 GEMM (ReLU fused in the epilogue)."""

    def __init__(self, in_features: int = 1024, hidden: int = 4096, depth: int = 4, out_features: int = 8) -> None:
        super().__init__()
        dims = [in_features] + [hidden] * depth
        layers = [bnn.Linear(a, b, relu=True, out_fp32=False) for a, b in zip(dims[:-1], dims[1:])]
        layers.append(bnn.Linear(hidden, out_features, out_fp32=True))
        self.layers = nn.Sequential(*layers)

    def forward(self, x):
        return self.layers(x)
