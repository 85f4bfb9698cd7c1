"""Single-process data parallelism replayed from CUDA graphs.

``DataParallel`` re-creates functional replicas and issues every replica's autograd functions from
Python on every step; beyond two GPUs the step time is the host's (measured on 8 x B200: 11.1 ms per
toy-MLP step, of which ~2 ms is GPU work).  ``GraphedDataParallel`` keeps PERSISTENT replicas whose
parameters and fp32 buffers are views of one flat "state" buffer per device and captures, per
device, ``zero grads -> forward -> loss -> backward`` ONCE as a CUDA graph.  A training step is then

    state broadcast (root flat -> replica flats, scatter + all-gather over NVSwitch)
    -> input scatter -> one graph launch per device
    -> gradient reduce-scatter (pull kernel per device) + gather of the reduced slices on the root,

i.e. a handful of launches and peer copies whatever the model size; the optimizer runs on the root's
parameters as with ``DataParallel`` (/root/reference/01.data_parallel.ipynb:275-287,478-487 is the
workflow this accelerates).  Semantics match ``DataParallel`` with a mean-reduced loss: the root
holds the model, the other devices only ever see copies, gradients arrive summed on the root.
"""
from __future__ import annotations

import copy
from typing import Callable, List, Sequence

import torch
from torch import nn

from .. import _lib
from .comm import _dev, _slices, _two_level_broadcast, ensure_peer_access


class GraphedDataParallel:
    """``module`` lives on ``device_ids[0]``; ``example_x`` / ``example_y`` fix the mini-batch shape
    (dim 0 is split evenly over the devices).  ``train_step(x, y)`` returns the mean loss (on the
    root) and leaves the summed gradients in the root parameters' ``.grad``; call your optimizer's
    ``step()`` afterwards -- the next ``train_step`` re-broadcasts the updated state."""

    def __init__(self, module: nn.Module, device_ids: Sequence, loss_fn: Callable,
                 example_x: torch.Tensor, example_y: torch.Tensor) -> None:
        self.devices = [_dev(d) for d in device_ids]
        if not self.devices or any(d.type != "cuda" for d in self.devices):
            raise ValueError("GraphedDataParallel needs CUDA devices (use DataParallel on the CPU)")
        n = len(self.devices)
        B = example_x.shape[0]
        if B % n:
            raise ValueError(f"batch {B} is not divisible over {n} devices")
        self.module, self.loss_fn = module, loss_fn
        self.batch, self.per = B, B // n
        self._x_shape, self._y_shape = tuple(example_x.shape), tuple(example_y.shape)
        root = self.devices[0]
        for t in list(module.parameters()) + list(module.buffers()):
            if t.device != root:
                raise RuntimeError(f"module must live on {root} (device_ids[0]), found a tensor on {t.device}")
        if n > 1:
            ensure_peer_access(self.devices)
        self.replicas: List[nn.Module] = [module] + [copy.deepcopy(module).to(d) for d in self.devices[1:]]
        self.streams = [torch.cuda.Stream(d) for d in self.devices]
        self._flatten_state()
        self._capture(example_x, example_y)

    # ------------------------------------------------------------------ flat state / gradients
    @staticmethod
    def _state_tensors(m: nn.Module):
        ps = [p for p in m.parameters()]
        bs = [b for b in m.buffers() if b.dtype == torch.float32]
        return ps, bs

    def _flatten_state(self) -> None:
        """Parameters and fp32 buffers of every replica become views of one flat buffer per device
        (same layout everywhere), gradients views of a second one: the broadcast and the reduction
        then move two contiguous ranges instead of hundreds of tensors."""
        self._state, self._grad = [], []
        self._params: List[List[torch.Tensor]] = []
        for m, dev in zip(self.replicas, self.devices):
            ps, bs = self._state_tensors(m)
            for t in ps + bs:
                if t.dtype != torch.float32:
                    raise ValueError("GraphedDataParallel expects fp32 parameters (bf16 operands are derived per step)")
            offs, total = [], 0
            for t in ps + bs:
                offs.append(total)
                total += (t.numel() + 63) // 64 * 64
            with torch.cuda.device(dev):
                flat = torch.zeros(total, device=dev, dtype=torch.float32)
                g_total = offs[len(ps)] if bs else total
                gflat = torch.zeros(max(g_total, 64), device=dev, dtype=torch.float32)
                with torch.no_grad():
                    for t, o in zip(ps + bs, offs):
                        view = flat[o:o + t.numel()].view(t.shape)
                        view.copy_(t)
                        t.data = view
                for p, o in zip(ps, offs):
                    p.grad = gflat[o:o + p.numel()].view(p.shape) if p.requires_grad else None
            self._state.append(flat)
            self._grad.append(gflat)
            self._params.append(ps)
        self._grad_views = [[p.grad for p in ps] for ps in self._params]
        n = len(self.devices)
        self._gslices = _slices(self._grad[0].numel(), n)
        self._red = []
        for d, dev in enumerate(self.devices):
            lo, hi = self._gslices[d]
            with torch.cuda.device(dev):
                self._red.append(torch.empty(max(hi - lo, 1), device=dev, dtype=torch.float32))

    def _attach_grads(self) -> None:
        """``optimizer.zero_grad(set_to_none=True)`` detaches the root's gradients; the captured
        graphs write the flat buffer, so point ``.grad`` back at it (no zeroing needed: every replay
        starts by clearing the buffer)."""
        for p, g in zip(self._params[0], self._grad_views[0]):
            if g is not None and p.grad is not g:
                p.grad = g

    def zero_grad(self) -> None:
        self._attach_grads()

    # ------------------------------------------------------------------ capture
    def _step_fn(self, d: int):
        self._grad[d].zero_()
        out = self.replicas[d](self.sx[d])
        loss = self.loss_fn(out, self.sy[d]) * (1.0 / len(self.devices))
        loss.backward()
        return loss.detach()

    def _capture(self, ex_x: torch.Tensor, ex_y: torch.Tensor) -> None:
        n, per = len(self.devices), self.per
        self.sx, self.sy = [], []
        for d, dev in enumerate(self.devices):
            with torch.cuda.device(dev):
                self.sx.append(ex_x[d * per:(d + 1) * per].to(dev, copy=True))
                self.sy.append(ex_y[d * per:(d + 1) * per].to(dev, copy=True))
        for dev in self.devices:
            torch.cuda.synchronize(dev)
        self.graphs, self.losses = [], []
        self.ev_in = [torch.cuda.Event() for _ in range(n)]
        self.ev_bwd = [torch.cuda.Event() for _ in range(n)]
        self.ev_red = [torch.cuda.Event() for _ in range(n)]
        self.ev_done = torch.cuda.Event()
        for d, dev in enumerate(self.devices):
            m = self.replicas[d]
            saved = [b.detach().clone() for b in m.buffers()]
            with torch.cuda.device(dev):
                with torch.cuda.stream(self.streams[d]):
                    self._step_fn(d)                      # lazy kernel setup, accumulator streams
                    self._step_fn(d)
                self.streams[d].synchronize()
                with torch.no_grad():
                    for b, s in zip(m.buffers(), saved):
                        b.copy_(s)
                for p in self._params[d]:                 # operand casts must be captured, not cached
                    if hasattr(p, "_b200_shadow"):
                        p._b200_shadow = None
                torch.cuda.synchronize(dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.streams[d]):
                    loss = self._step_fn(d)
                self.graphs.append(g)
                self.losses.append(loss)
                self._grad[d].zero_()
                torch.cuda.synchronize(dev)
        self._attach_grads()

    # ------------------------------------------------------------------ one training step
    def _broadcast_state(self) -> None:
        root, flats = self.devices[0], self._state
        if len(flats) == 1:
            return
        if len(flats) >= 3:
            _two_level_broadcast(flats[0], flats[1:], root)
        else:
            flats[1].copy_(flats[0], non_blocking=True)
        # the replicas' graphs run on their own streams: order them after the copies just queued
        for d in range(1, len(flats)):
            ev = self.ev_in[d]
            ev.record(torch.cuda.current_stream(self.devices[d]))
            self.streams[d].wait_event(ev)

    def _reduce_grads(self) -> None:
        n = len(self.devices)
        if n == 1:
            return
        L = _lib.lib()
        for d in range(n):
            self.ev_bwd[d].record(self.streams[d])
        srcs = self._grad
        for d in range(n):
            lo, hi = self._gslices[d]
            if hi <= lo:
                continue
            s = self.streams[d]
            for k in range(n):
                if k != d:
                    s.wait_event(self.ev_bwd[k])
            with torch.cuda.device(self.devices[d]):
                _lib.check(L.b200dt_p2p_reduce_f32(self._red[d].data_ptr(),
                                                   _lib.u64_array([f.data_ptr() + lo * 4 for f in srcs]),
                                                   n, hi - lo, 1.0, 0, s.cuda_stream), "p2p_reduce")
            self.ev_red[d].record(s)
        root_s = self.streams[0]
        for d in range(n):
            lo, hi = self._gslices[d]
            if hi <= lo:
                continue
            root_s.wait_event(self.ev_red[d])
            # both ends of the peer copy are pinned to OUR streams (torch runs it on the source's current one)
            with torch.cuda.stream(self.streams[d]), torch.cuda.stream(root_s):
                self._grad[0][lo:hi].copy_(self._red[d][:hi - lo], non_blocking=True)

    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if tuple(x.shape) != self._x_shape or tuple(y.shape) != self._y_shape:
            raise ValueError(f"GraphedDataParallel was captured for x{self._x_shape} / y{self._y_shape}, "
                             f"got x{tuple(x.shape)} / y{tuple(y.shape)}")
        n, per = len(self.devices), self.per
        self._attach_grads()
        cur = [torch.cuda.current_stream(d) for d in self.devices]
        for s, c in zip(self.streams, cur):
            s.wait_stream(c)
            s.wait_event(self.ev_done)           # the previous step's reduction has read every gradient buffer
        # the updated parameters travel from the root's flat state to the replicas'
        with torch.cuda.stream(self.streams[0]):
            self._broadcast_state()
        for d in range(n):
            with torch.cuda.stream(self.streams[d]):
                self.sx[d].copy_(x[d * per:(d + 1) * per], non_blocking=True)
                self.sy[d].copy_(y[d * per:(d + 1) * per], non_blocking=True)
                self.graphs[d].replay()
        self._reduce_grads()
        root_s = self.streams[0]
        with torch.cuda.stream(root_s):
            parts = [self.losses[0]]
            for d in range(1, n):
                root_s.wait_event(self.ev_bwd[d])
                with torch.cuda.stream(self.streams[d]):
                    parts.append(self.losses[d].to(self.devices[0], non_blocking=True))
            loss = torch.stack(parts).sum() if n > 1 else parts[0].clone()
            self.ev_done.record(root_s)
        for s, c in zip(self.streams, cur):
            c.wait_stream(s)
        return loss

    @property
    def launches_per_step(self) -> int:
        """Graph launches per ``train_step`` (peer copies and the pull-reduce kernels come on top)."""
        return len(self.devices)
