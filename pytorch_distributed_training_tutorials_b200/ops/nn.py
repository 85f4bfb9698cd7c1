"""``nn.Module`` wrappers over the differentiable sm_100a ops.

Parameter names and ``state_dict`` layouts match ``torch.nn`` / torchvision (conv weights are
exported as ``[Cout, Cin, kH, kW]``) so checkpoints round-trip with plain PyTorch models; inside
the module convolution filters live as ``[Cout, kH, kW, Cin]`` fp32, the layout the TMA maps read.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from . import functional as Fn


class Linear(nn.Module):
    """``nn.Linear`` on the tcgen05 GEMM (bf16 operands, fp32 accumulate, fp32 master weights)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, relu: bool = False,
                 out_fp32: bool = True) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.relu, self.out_fp32 = relu, out_fp32
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        self.reset_parameters()

    def reset_parameters(self) -> None:
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.in_features) if self.in_features > 0 else 0
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            y = torch.nn.functional.linear(x, self.weight, self.bias)
            return torch.relu(y) if self.relu else y
        if self.skinny:
            # feature counts the TMA / tcgen05 path cannot take: fp32 SIMT kernel, no library GEMM
            y = Fn.skinny_linear(x, self.weight, self.bias)
            return torch.relu(y) if self.relu else y
        return Fn.linear(x, self.weight, self.bias, self.relu, self.out_fp32)

    @property
    def skinny(self) -> bool:
        """True for the degenerate layers (K or N not a multiple of 8 and small): the tutorial's
        Linear(20,1), (32,2), (10000,10), (10,5)."""
        return self.in_features % 8 != 0 or (self.out_features % 8 != 0
                                             and self.in_features * self.out_features <= (1 << 20))

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"


class Conv2d(nn.Module):
    """Bias-free NHWC convolution, kernel 1 or 3, padding k//2, stride 1 or 2."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, stride: int = 1) -> None:
        super().__init__()
        assert kernel_size in (1, 3) and stride in (1, 2)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride = kernel_size, stride
        w = torch.empty(out_channels, in_channels, kernel_size, kernel_size)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w.permute(0, 2, 3, 1).contiguous())     # OHWI
        self._register_state_dict_hook(_export_oihw)
        self._register_load_state_dict_pre_hook(_import_oihw, with_module=True)

    def forward(self, x: torch.Tensor, with_stats: bool = False):
        """``with_stats=True`` -> ``(y, sums)`` for a following ``BatchNorm2d(y, sums=sums)``."""
        if not x.is_cuda:
            # CPU reference path (plain torch, same NHWC interface): debugging and the CPU test-suite
            y = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), self.weight.permute(0, 3, 1, 2).to(x.dtype),
                                           stride=self.stride, padding=self.kernel_size // 2).permute(0, 2, 3, 1)
            return (y, None) if with_stats else y
        return Fn.conv2d(x, self.weight, self.stride, with_stats)

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, layout=NHWC")


class StemConv(nn.Module):
    """ResNet stem: 7x7 stride-2 pad-3 conv from the fp32 NCHW input to NHWC bf16."""

    def __init__(self, out_channels: int = 64) -> None:
        super().__init__()
        w = torch.empty(out_channels, 3, 7, 7)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w.permute(0, 2, 3, 1).contiguous())
        self._register_state_dict_hook(_export_oihw)
        self._register_load_state_dict_pre_hook(_import_oihw, with_module=True)

    def forward(self, x: torch.Tensor, with_stats: bool = False):
        if not x.is_cuda:
            y = torch.nn.functional.conv2d(x, self.weight.permute(0, 3, 1, 2).to(x.dtype), stride=2,
                                           padding=3).permute(0, 2, 3, 1)
            return (y, None) if with_stats else y
        return Fn.stem_conv(x, self.weight, with_stats)


def _export_oihw(module, state_dict, prefix, local_metadata):
    key = prefix + "weight"
    if key in state_dict:
        state_dict[key] = state_dict[key].permute(0, 3, 1, 2).contiguous()
    return state_dict


def _import_oihw(module, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
    # checkpoints always carry torch's [Cout, Cin, kH, kW]; the module keeps [Cout, kH, kW, Cin]
    key = prefix + "weight"
    if key in state_dict and state_dict[key].dim() == 4:
        state_dict[key] = state_dict[key].permute(0, 2, 3, 1).contiguous()


class BatchNorm2d(nn.Module):
    """Training-mode BatchNorm over NHWC bf16, fused with ReLU and an optional residual add."""

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1, relu: bool = False) -> None:
        super().__init__()
        self.num_features, self.eps, self.momentum, self.relu = num_features, eps, momentum, relu
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None,
                sums: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not x.is_cuda:
            # CPU reference path: torch's batch_norm on the NCHW view, then the fused tail in plain ops
            if self.training:
                self.num_batches_tracked += 1
            y = torch.nn.functional.batch_norm(x.permute(0, 3, 1, 2), self.running_mean, self.running_var,
                                               self.weight, self.bias, self.training, self.momentum,
                                               self.eps).permute(0, 2, 3, 1)
            if residual is not None:
                y = y + residual
            return torch.relu(y) if self.relu else y
        # the step counter is bumped by the apply kernel itself (53 fewer launches per ResNet-50 step)
        return Fn.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, residual,
                             self.eps, self.momentum, self.relu, self.training, sums,
                             self.num_batches_tracked if self.training else None)

    def extra_repr(self) -> str:
        return f"{self.num_features}, eps={self.eps}, momentum={self.momentum}, relu={self.relu}"


class MaxPool3x3s2(nn.Module):
    def forward(self, x):
        if not x.is_cuda:
            return torch.nn.functional.max_pool2d(x.permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)
        return Fn.max_pool_3x3_s2(x)


class GlobalAvgPool(nn.Module):
    def forward(self, x):
        if not x.is_cuda:
            return x.mean((1, 2))
        return Fn.global_avg_pool(x)


class MSELoss(nn.Module):
    def forward(self, pred, target):
        if not pred.is_cuda:
            return torch.nn.functional.mse_loss(pred, target)
        return Fn.mse_loss(pred, target)


class CrossEntropyLoss(nn.Module):
    def forward(self, logits, target):
        if not logits.is_cuda:
            return torch.nn.functional.cross_entropy(logits, target)
        return Fn.cross_entropy(logits, target)
