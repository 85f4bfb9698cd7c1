"""Pipelined model parallelism replayed from CUDA graphs: one Python thread, every GPU busy.

``Pipeline`` (model_parallel.py) issues each micro-batch of each stage as a few hundred eager kernel
launches; a B200 retires a 64-image ResNet-50 micro-batch faster than Python can issue it, so the
second GPU mostly waits for the host and pipelining buys ~10 % over the naive split.  Here every
(stage, micro-batch slot) pair is captured ONCE into two CUDA graphs -- forward (+ loss on the last
stage) and backward -- and a training step is ``2 * stages * chunks`` graph launches plus the peer
copies of the boundary activations / gradients between them, ordered with events.  The host cost of
a micro-batch drops from milliseconds to tens of microseconds and the fill-drain schedule becomes
GPU-bound: stage k runs micro-batch i while stage k+1 runs micro-batch i-1.

All slots of one stage share a graph memory pool and are captured in the order they replay
(forward 0..m-1, backward m-1..0), so the activations a slot saves for its backward stay untouched
until that backward has run -- the same memory a GPipe schedule holds live in eager mode.
Parameter gradients accumulate in place into one flat fp32 buffer per device (``p.grad`` are views),
which the captured backward graphs write directly; the optimizer step stays outside the graphs.

The reference has only the naive split (/root/reference/03.model_parallel.ipynb:807-833) and eager
execution; the pipelined variant is BASELINE.json config 4.
"""
from __future__ import annotations

from typing import Callable, Iterable, List, Optional, Sequence

import torch
from torch import nn

from .comm import _dev


def _owner_module(stage) -> Optional[nn.Module]:
    if isinstance(stage, nn.Module):
        return stage
    owner = getattr(stage, "__self__", None)
    return owner if isinstance(owner, nn.Module) else None


class GraphedPipeline:
    """``stages[k]`` (a module or any callable of one tensor) runs on ``devices[k]``; parameters must
    already live there.  ``example_x`` / ``example_y`` fix the mini-batch shape: they are split into
    ``chunks`` equal micro-batches.  ``train_step(x, y)`` runs forward + loss + backward of one
    mini-batch and returns the mean loss (a tensor on the last device); gradients are left in
    ``p.grad`` for any optimizer.  ``zero_grad()`` clears them with one memset per device."""

    def __init__(self, stages: Sequence[Callable], devices: Sequence, chunks: int, loss_fn: Callable,
                 example_x: torch.Tensor, example_y: torch.Tensor,
                 parameters: Optional[Iterable[torch.Tensor]] = None) -> None:
        assert len(stages) == len(devices) and len(stages) >= 1
        self.stages = list(stages)
        self.devices = [_dev(d) for d in devices]
        if any(d.type != "cuda" for d in self.devices):
            raise ValueError("GraphedPipeline needs CUDA devices (use Pipeline on the CPU)")
        self.loss_fn = loss_fn
        n = len(self.stages)
        B = example_x.shape[0]
        if B % chunks != 0:
            raise ValueError(f"batch {B} is not divisible into {chunks} equal micro-batches")
        self.chunks, self.batch = chunks, B
        self._x_shape, self._y_shape = tuple(example_x.shape), tuple(example_y.shape)
        if parameters is None:
            seen, plist = set(), []
            for st in self.stages:
                mod = _owner_module(st)
                if mod is None:
                    raise ValueError("pass parameters= when the stages are plain callables")
                for p in mod.parameters():
                    if id(p) not in seen:
                        seen.add(id(p)); plist.append(p)
            parameters = plist
        self.params: List[torch.Tensor] = [p for p in parameters if p.requires_grad]
        self._buffers: List[torch.Tensor] = []
        seen = set()
        for st in self.stages:
            mod = _owner_module(st)
            if mod is not None:
                for b in mod.buffers():
                    if id(b) not in seen:
                        seen.add(id(b)); self._buffers.append(b)
        self.streams = [torch.cuda.Stream(d) for d in self.devices]
        self.pools = [torch.cuda.graph_pool_handle() for _ in self.devices]
        self._make_flat_grads()
        self._capture(example_x, example_y)

    # ------------------------------------------------------------------ gradients
    def _make_flat_grads(self) -> None:
        by_dev = {}
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise ValueError("GraphedPipeline expects contiguous fp32 parameters")
            by_dev.setdefault(p.device, []).append(p)
        self._flat, self._grads = {}, {}
        for dev, ps in by_dev.items():
            offs, total = [], 0
            for p in ps:
                offs.append(total)
                total += (p.numel() + 63) // 64 * 64          # 256-byte aligned views
            flat = torch.zeros(total, device=dev, dtype=torch.float32)
            self._flat[dev] = flat
            for p, o in zip(ps, offs):
                self._grads[id(p)] = flat[o:o + p.numel()].view(p.shape)
        self._attach_grads(force=True)

    def _attach_grads(self, force: bool = False) -> None:
        """The captured backward graphs add into the flat buffers; make ``p.grad`` point at them again
        after an ``optimizer.zero_grad(set_to_none=True)``."""
        stale = force or any(p.grad is not self._grads[id(p)] for p in self.params)
        if not stale:
            return
        for dev, flat in self._flat.items():
            with torch.cuda.device(dev):
                flat.zero_()
        for p in self.params:
            p.grad = self._grads[id(p)]

    def zero_grad(self) -> None:
        self._attach_grads(force=True)

    # ------------------------------------------------------------------ capture
    def _capture(self, ex_x: torch.Tensor, ex_y: torch.Tensor) -> None:
        n, m = len(self.stages), self.chunks
        devs, streams = self.devices, self.streams
        mb = self.batch // m
        self._w = 1.0 / m
        # static boundary buffers (filled below once the shapes are known)
        self.sin: List[List[torch.Tensor]] = [[None] * m for _ in range(n)]
        self.out: List[List[torch.Tensor]] = [[None] * m for _ in range(n)]
        self.gout: List[List[torch.Tensor]] = [[None] * m for _ in range(n)]
        self.gin: List[List[torch.Tensor]] = [[None] * m for _ in range(n)]
        self.fwd = [[None] * m for _ in range(n)]
        self.bwd = [[None] * m for _ in range(n)]
        self.ev_f = [[torch.cuda.Event() for _ in range(m)] for _ in range(n)]
        self.ev_b = [[torch.cuda.Event() for _ in range(m)] for _ in range(n)]
        with torch.cuda.device(devs[0]):
            self.sin[0] = [ex_x[i * mb:(i + 1) * mb].to(devs[0], copy=True) for i in range(m)]
        with torch.cuda.device(devs[-1]):
            self.sy = [ex_y[i * mb:(i + 1) * mb].to(devs[-1], copy=True) for i in range(m)]
        for d in devs:
            torch.cuda.synchronize(d)

        # ---- one eager micro-batch through every stage on the side streams: lazy initialisation of
        # the kernels and the boundary shapes; parameters, gradients and buffers are restored after
        saved = [b.detach().clone() for b in self._buffers]
        shapes = []
        a = self.sin[0][0]
        acts = []
        for k in range(n):
            with torch.cuda.device(devs[k]), torch.cuda.stream(streams[k]):
                if k > 0:
                    a = a.detach().to(devs[k]).requires_grad_(a.is_floating_point())
                    shapes.append((tuple(a.shape), a.dtype))
                o = self.stages[k](a)
                if k == n - 1:
                    o = self.loss_fn(o, self.sy[0]) * self._w
                acts.append((a, o))
                streams[k].synchronize()
                a = o
        g = None
        for k in reversed(range(n)):
            a, o = acts[k]
            with torch.cuda.device(devs[k]), torch.cuda.stream(streams[k]):
                if k == n - 1:
                    o.backward()
                else:
                    o.backward(g.to(devs[k]))
                streams[k].synchronize()
                g = a.grad if k > 0 else None
        del acts, a, o, g
        with torch.no_grad():
            for b, s in zip(self._buffers, saved):
                b.copy_(s)
        del saved
        self._attach_grads(force=True)
        for p in self.params:                     # the operand casts must be IN the graphs (slot 0)
            if hasattr(p, "_b200_shadow"):
                p._b200_shadow = None
        for k in range(1, n):
            shp, dt = shapes[k - 1]
            with torch.cuda.device(devs[k]):
                self.sin[k] = [torch.zeros(shp, dtype=dt, device=devs[k]).requires_grad_(dt.is_floating_point)
                               for _ in range(m)]
            with torch.cuda.device(devs[k - 1]):
                self.gout[k - 1] = [torch.zeros(shp, dtype=dt, device=devs[k - 1]) for _ in range(m)]
        for d in devs:
            torch.cuda.synchronize(d)

        # ---- forward graphs, in replay order per stage
        live = [[None] * m for _ in range(n)]
        for k in range(n):
            with torch.cuda.device(devs[k]):
                for i in range(m):
                    gph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gph, pool=self.pools[k], stream=streams[k]):
                        o = self.stages[k](self.sin[k][i])
                        if k == n - 1:
                            o = self.loss_fn(o, self.sy[i]) * self._w
                    self.fwd[k][i] = gph
                    live[k][i] = o
                    self.out[k][i] = o.detach()
        # ---- backward graphs, reverse order
        for k in reversed(range(n)):
            with torch.cuda.device(devs[k]):
                for i in reversed(range(m)):
                    gph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gph, pool=self.pools[k], stream=streams[k]):
                        if k == n - 1:
                            live[k][i].backward()
                        else:
                            live[k][i].backward(self.gout[k][i])
                    self.bwd[k][i] = gph
                    live[k][i] = None
                    if k > 0:
                        self.gin[k][i] = self.sin[k][i].grad
        for d in devs:
            torch.cuda.synchronize(d)
        # capture records but does not run: nothing was accumulated, the flat buffers are still zero

    # ------------------------------------------------------------------ replay
    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if tuple(x.shape) != self._x_shape or tuple(y.shape) != self._y_shape:
            raise ValueError(f"GraphedPipeline was captured for x{self._x_shape} / y{self._y_shape}, "
                             f"got x{tuple(x.shape)} / y{tuple(y.shape)}")
        n, m = len(self.stages), self.chunks
        mb = self.batch // m
        streams = self.streams
        self._attach_grads()
        cur = [torch.cuda.current_stream(d) for d in self.devices]
        for s, c in zip(streams, cur):
            s.wait_stream(c)
        with torch.cuda.stream(streams[0]):
            for i in range(m):
                self.sin[0][i].copy_(x[i * mb:(i + 1) * mb], non_blocking=True)
        with torch.cuda.stream(streams[-1]):
            for i in range(m):
                self.sy[i].copy_(y[i * mb:(i + 1) * mb], non_blocking=True)
        for i in range(m):
            for k in range(n):
                s = streams[k]
                with torch.cuda.stream(s):
                    if k > 0:
                        s.wait_event(self.ev_f[k - 1][i])
                        self.sin[k][i].detach().copy_(self.out[k - 1][i], non_blocking=True)
                    self.fwd[k][i].replay()
                    if k < n - 1:
                        self.ev_f[k][i].record(s)
        for i in reversed(range(m)):
            for k in reversed(range(n)):
                s = streams[k]
                with torch.cuda.stream(s):
                    if k < n - 1:
                        s.wait_event(self.ev_b[k + 1][i])
                        self.gout[k][i].copy_(self.gin[k + 1][i], non_blocking=True)
                    self.bwd[k][i].replay()
                    if k > 0:
                        self.ev_b[k][i].record(s)
        with torch.cuda.stream(streams[-1]):
            loss = torch.stack(self.out[-1]).sum() if m > 1 else self.out[-1][0].clone()
        for s, c in zip(streams, cur):
            c.wait_stream(s)
        return loss

    @property
    def launches_per_step(self) -> int:
        """Graph launches per ``train_step`` (the boundary copies and the loss sum come on top)."""
        return 2 * len(self.stages) * self.chunks
