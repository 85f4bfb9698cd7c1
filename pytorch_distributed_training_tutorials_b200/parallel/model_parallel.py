"""Naive (one micro-batch) and pipelined (GPipe-style) model parallelism inside one process.

``StagedModule`` pins consecutive stages of a model to different devices and moves the
activation between them, which is what the reference does by hand at
/root/reference/03.model_parallel.ipynb:440-450 (toy) and :807-833 (ResNet-50 split after
``layer2``): only one GPU is busy at a time.  ``Pipeline`` splits the batch into micro-batches
and issues stage k on micro-batch i while stage k+1 works on micro-batch i-1; the boundary
activation travels on a dedicated copy stream so the hop overlaps compute on both sides.  The
pipelined variant is not in the reference; it is required by BASELINE.json (config 4).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
from torch import nn

from .comm import _dev

_copy_streams = {}


def _copy_stream(dev: torch.device) -> torch.cuda.Stream:
    s = _copy_streams.get(dev)
    if s is None:
        s = torch.cuda.Stream(dev)
        _copy_streams[dev] = s
    return s


def _hop_raw(x: torch.Tensor, dst: torch.device) -> torch.Tensor:
    """Peer copy on the source device's copy stream; ``dst``'s compute stream waits for it."""
    src = x.device
    if src == dst:
        return x
    if src.type != "cuda" or dst.type != "cuda":
        return x.to(dst)
    cs = _copy_stream(src)
    done_compute = torch.cuda.Event()
    done_compute.record(torch.cuda.current_stream(src))
    cs.wait_event(done_compute)
    with torch.cuda.device(dst):
        out = torch.empty_like(x, device=dst)
    ready = torch.cuda.Event()
    ready.record(torch.cuda.current_stream(dst))      # allocation-order safety on dst
    cs.wait_event(ready)
    with torch.cuda.stream(cs):
        out.copy_(x, non_blocking=True)
    x.record_stream(cs)
    out.record_stream(cs)
    copied = torch.cuda.Event()
    copied.record(cs)
    torch.cuda.current_stream(dst).wait_event(copied)
    return out


class _Hop(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dst):
        ctx.src = x.device
        return _hop_raw(x.contiguous(), dst)

    @staticmethod
    def backward(ctx, g):
        return _hop_raw(g.contiguous(), ctx.src), None


def hop(x: torch.Tensor, device) -> torch.Tensor:
    """Differentiable stage-boundary transfer (gradient hops back the same way)."""
    device = _dev(device)
    if x.device == device:
        return x
    return _Hop.apply(x, device)


class StagedModule(nn.Module):
    """Stages on fixed devices, one activation hop per boundary (naive model parallelism)."""

    def __init__(self, stages: Sequence[nn.Module], devices: Sequence) -> None:
        super().__init__()
        assert len(stages) == len(devices) and len(stages) >= 1
        self.devices = [_dev(d) for d in devices]
        self.stages = nn.ModuleList([s.to(d) for s, d in zip(stages, self.devices)])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for stage, dev in zip(self.stages, self.devices):
            x = stage(hop(x, dev))
        return x


class Pipeline(nn.Module):
    """Micro-batched execution of ``StagedModule``-style stages (fill-drain schedule).

    At clock tick t stage k runs micro-batch t-k; stages are issued last-to-first within a tick so
    that a stage's input hop is already queued when the producer moves on.  All launches are
    asynchronous, so one Python thread keeps every device busy; autograd replays the schedule in
    reverse for backward.  ``split_size`` is the micro-batch size."""

    def __init__(self, stages: Sequence[nn.Module], devices: Sequence, split_size: int) -> None:
        super().__init__()
        assert split_size >= 1
        self.devices = [_dev(d) for d in devices]
        self.stages = nn.ModuleList([s.to(d) for s, d in zip(stages, self.devices)])
        self.split_size = split_size

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        micro = list(x.split(self.split_size, dim=0))
        m, n = len(micro), len(self.stages)
        acts: List[Optional[torch.Tensor]] = list(micro)
        outs: List[Optional[torch.Tensor]] = [None] * m
        for t in range(m + n - 1):
            for k in reversed(range(n)):
                i = t - k
                if 0 <= i < m:
                    y = self.stages[k](hop(acts[i], self.devices[k]))
                    if k == n - 1:
                        outs[i] = y
                        acts[i] = None
                    else:
                        acts[i] = y
        return torch.cat(outs, dim=0)


def split_sequential(module: nn.Sequential, boundaries: Sequence[int]) -> List[nn.Sequential]:
    """Cut an ``nn.Sequential`` before the given child indices: ``[2, 5]`` -> [0:2], [2:5], [5:]."""
    kids = list(module.children())
    cuts = [0] + list(boundaries) + [len(kids)]
    return [nn.Sequential(*kids[a:b]) for a, b in zip(cuts[:-1], cuts[1:])]


def model_size(model: nn.Module) -> int:
    """Number of parameters, de-duplicated (reference: 03.model_parallel.ipynb:844-845)."""
    return sum(p.numel() for p in model.parameters())
