"""Whole-step CUDA graph capture.

A training step of this library is several hundred small-to-medium kernel launches driven from
Python; once shapes are static the whole forward + loss + backward + gradient all-reduce +
optimizer update is captured ONCE into a CUDA graph and replayed with a single launch per step,
so the host stops being the bottleneck for small batches / short kernels.  (The reference runs
eager PyTorch only; this is the B200-side replacement for a tracing compiler: streams + graphs.)

Works with: the sm_100a ops of this package, ``FusedSGD``, ``DistributedDataParallel`` (the bucket
kernels run on the side stream inside the captured region; every rank must capture and replay in
lock-step), ``FusedAdam(capturable=True)`` (device-side step count).  Not capturable: data-dependent
Python control flow, optimizers that keep their step count on the host.
"""
from __future__ import annotations

from typing import Callable

import torch

from .. import _lib


class GraphedTrainStep:
    """``step(x, y) -> loss`` replaying a captured ``zero_grad / forward / loss / backward / step``."""

    def __init__(self, model: torch.nn.Module, loss_fn: Callable, optimizer: torch.optim.Optimizer,
                 example_x: torch.Tensor, example_y: torch.Tensor, warmup: int = 3) -> None:
        if not example_x.is_cuda:
            raise ValueError("GraphedTrainStep needs CUDA tensors")
        self.model, self.loss_fn, self.optimizer = model, loss_fn, optimizer
        self.device = example_x.device
        self.static_x = example_x.clone()
        self.static_y = example_y.clone()
        self.kernels_per_step = 0
        cur = torch.cuda.current_stream(self.device)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            # gradient accumulators must not be bound to the legacy default stream (see the reducer)
            for m in model.modules():
                if hasattr(m, "rebind_autograd_hooks"):
                    m.rebind_autograd_hooks()
            # the warm-up iterations (lazy kernel setup; DDP needs >= 2 to settle its buckets) must not
            # train the model: parameters, buffers and optimizer state are put back afterwards
            snap = self._snapshot()
            for _ in range(max(warmup, 3)):
                self._eager()
            self._restore(snap)
            del snap
        cur.wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        self.optimizer.zero_grad(set_to_none=True)
        before = _lib.launch_count + self._reducer_launches()
        with torch.cuda.graph(self.graph):
            self.static_loss = self._eager(zero=False)
        self.kernels_per_step = _lib.launch_count + self._reducer_launches() - before

    def _reducer_launches(self) -> int:
        """Collectives the C++ reducer launched itself (they never pass through the ctypes counter)."""
        n = 0
        for m in self.model.modules():
            r = getattr(m, "_reducer", None)
            if r is not None and hasattr(r, "stats"):
                n += int(r.stats().get("native_launches", 0))
        return n

    def _snapshot(self):
        import copy
        tensors = list(self.model.parameters()) + list(self.model.buffers())
        return ([t.detach().clone() for t in tensors], tensors, copy.deepcopy(self.optimizer.state_dict()))

    @torch.no_grad()
    def _restore(self, snap) -> None:
        saved, tensors, opt_state = snap
        for t, s in zip(tensors, saved):
            t.copy_(s)
        self.optimizer.load_state_dict(opt_state)
        if hasattr(self.optimizer, "refresh_shadows"):
            self.optimizer.refresh_shadows()

    def _eager(self, zero: bool = True):
        if zero:
            self.optimizer.zero_grad(set_to_none=True)
        loss = self.loss_fn(self.model(self.static_x), self.static_y)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def matches(self, x: torch.Tensor, y: torch.Tensor) -> bool:
        return (x.shape == self.static_x.shape and y.shape == self.static_y.shape
                and x.dtype == self.static_x.dtype and y.dtype == self.static_y.dtype)

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if x.data_ptr() != self.static_x.data_ptr():
            self.static_x.copy_(x, non_blocking=True)
        if y.data_ptr() != self.static_y.data_ptr():
            self.static_y.copy_(y, non_blocking=True)
        self.graph.replay()
        _lib.launch_count += self.kernels_per_step      # replayed kernels of this library
        return self.static_loss
