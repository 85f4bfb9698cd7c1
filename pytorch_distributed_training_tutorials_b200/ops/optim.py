"""Fused multi-tensor optimizers.

``FusedSGD`` / ``FusedAdam`` are drop-in ``torch.optim.Optimizer`` subclasses with the hyper-
parameters the reference uses (SGD without momentum: /root/reference/ddp_gpus.py:82,
03.model_parallel.ipynb:534,977; Adam defaults: 01.data_parallel.ipynb:287).  On CUDA one kernel
launch per device updates every parameter, refreshes the bf16 operand shadows used by the tcgen05
kernels and can clear the gradients in the same pass; parameters may live on several devices
(the model-parallel case: one optimizer over ``cuda:0`` and ``cuda:1`` parameters).
On CPU tensors the same math runs through plain torch ops.
"""
from __future__ import annotations

from typing import Dict, List

import torch
from torch.optim import Optimizer

from .. import _lib
from .._lib import check, lib, ptr, stream_ptr

CHUNK = 8192


class _DeviceTable:
    """Device-resident {param, grad, shadow, numel, m, v} rows + the chunk list for one device."""

    def __init__(self, device: torch.device) -> None:
        self.device = device
        self.key = None
        self.fast_key = None          # (param ptrs..., grad ptrs...) of the last update: see _FusedBase._table
        self.meta = None
        self.chunks = None
        self.n_chunks = 0
        self._uploader = _lib.PinnedUploader()

    def update(self, rows: List[List[int]]) -> None:
        key = tuple(tuple(r) for r in rows)
        if key == self.key:
            return
        ch = []
        for ti, r in enumerate(rows):
            n = r[3]
            ch.extend((ti, s) for s in range(0, n, CHUNK))
        # stream-ordered uploads from library-owned pinned staging (capturable: they become memcpy
        # nodes of the graph, and torch's pinned allocator never sees them)
        meta_host = torch.tensor(rows, dtype=torch.int64)
        chunks_host = torch.tensor(ch, dtype=torch.int32)
        if self.meta is None or self.meta.numel() != meta_host.numel():
            self.meta = torch.empty(meta_host.shape, dtype=torch.int64, device=self.device)
        if self.chunks is None or self.chunks.numel() != chunks_host.numel():
            self.chunks = torch.empty(chunks_host.shape, dtype=torch.int32, device=self.device)
        stream = torch.cuda.current_stream(self.device)
        self._uploader.upload(self.meta, meta_host, stream)
        self._uploader.upload(self.chunks, chunks_host, stream)
        self.n_chunks = len(ch)
        self.key = key


class _FusedBase(Optimizer):
    def __init__(self, params, defaults, fuse_zero_grad: bool, bf16_shadow: bool):
        super().__init__(params, defaults)
        self.fuse_zero_grad = fuse_zero_grad
        self.bf16_shadow = bf16_shadow
        self._tables: Dict[tuple, _DeviceTable] = {}
        self._shadows: Dict[int, torch.Tensor] = {}     # kept out of .state: state_dict stays torch-compatible
        self._grads_cleared = False

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self.fuse_zero_grad and self._grads_cleared:
            # the last step() already zeroed every gradient in place
            self._grads_cleared = False
            return
        super().zero_grad(set_to_none=set_to_none)

    def _shadow_for(self, p: torch.Tensor):
        if not self.bf16_shadow or p.dim() < 2:
            return None
        sh = self._shadows.get(id(p))
        if sh is None or sh.device != p.device or sh.shape != p.shape:
            from . import native as NV
            sh = NV.cast_bf16(p.detach().contiguous())
            self._shadows[id(p)] = sh
        return sh

    def _cuda_groups(self, group):
        """params with grads of this param group, split by device."""
        by_dev: Dict[torch.device, List[torch.Tensor]] = {}
        cpu: List[torch.Tensor] = []
        for p in group["params"]:
            if p.grad is None:
                continue
            if p.grad.is_sparse:
                raise RuntimeError("sparse gradients are not supported")
            if p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.grad.is_contiguous() \
                    and p.grad.dtype == torch.float32:
                by_dev.setdefault(p.device, []).append(p)
            else:
                cpu.append(p)
        return by_dev, cpu

    def _table(self, gi: int, dev: torch.device, ps, build_rows) -> "_DeviceTable":
        """The device table of (param group, device), rebuilt only when a parameter or gradient address changed.
        In steady state the caching allocator hands every gradient the address it had a step ago, so a step costs
        2 x len(ps) ``data_ptr()`` calls and one tuple compare instead of building and hashing the row lists
        (1.2 ms of an 11 ms host-bound ResNet-50 step).  Optimizer state tensors (bf16 shadows, Adam moments) are
        created once and keep their addresses; ``load_state_dict`` drops the fast keys."""
        tab = self._tables.get((gi, dev))
        if tab is None:
            tab = self._tables[(gi, dev)] = _DeviceTable(dev)
        fast = tuple([p.data_ptr() for p in ps] + [p.grad.data_ptr() for p in ps])
        if tab.fast_key != fast:
            tab.update(build_rows())
            tab.fast_key = fast
        return tab

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        for tab in self._tables.values():
            tab.fast_key = None

    def refresh_shadows(self) -> None:
        """Re-derive the bf16 operand copies from the fp32 masters (after the masters were written by
        something other than ``step``: a checkpoint load, a restore after warm-up iterations)."""
        from . import native as NV
        for group in self.param_groups:
            for p in group["params"]:
                sh = self._shadows.get(id(p))
                if sh is not None and p.is_cuda and sh.device == p.device and sh.shape == p.shape:
                    NV.cast_bf16(p.detach().contiguous(), out=sh)
                    p._b200_shadow = (p._version, sh)

    def _publish_shadows(self, params) -> None:
        # the kernel wrote the parameters behind autograd's back: bump their version counters (what an in-place
        # torch op would have done) so that version-keyed caches -- DataParallel's replica cache, the bf16 operand
        # shadows -- notice, then re-publish the shadows the kernel has just refreshed under the new version
        torch.autograd.graph.increment_version(list(params))
        for p in params:
            sh = self._shadows.get(id(p))
            if sh is not None:
                p._b200_shadow = (p._version, sh)


class FusedSGD(_FusedBase):
    """SGD without momentum (optionally with weight decay): ``p -= lr * (g + wd * p)``."""

    def __init__(self, params, lr: float = 1e-3, weight_decay: float = 0.0, fuse_zero_grad: bool = False,
                 bf16_shadow: bool = True) -> None:
        if lr < 0:
            raise ValueError(f"Invalid learning rate: {lr}")
        super().__init__(params, dict(lr=lr, weight_decay=weight_decay), fuse_zero_grad, bf16_shadow)

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            by_dev, slow = self._cuda_groups(group)
            for dev, ps in by_dev.items():
                def rows(ps=ps):
                    out = []
                    for p in ps:
                        sh = self._shadow_for(p)
                        out.append([p.data_ptr(), p.grad.data_ptr(), sh.data_ptr() if sh is not None else 0,
                                    p.numel(), 0, 0, 0, 0])
                    return out
                tab = self._table(gi, dev, ps, rows)
                with torch.cuda.device(dev):
                    check(lib().b200dt_multi_sgd(ptr(tab.meta), ptr(tab.chunks), tab.n_chunks, group["lr"],
                                                 group["weight_decay"], grad_scale, int(self.fuse_zero_grad),
                                                 stream_ptr(dev)), "multi_sgd")
                self._publish_shadows(ps)
            for p in slow:
                g = p.grad * grad_scale if grad_scale != 1.0 else p.grad
                if group["weight_decay"]:
                    g = g.add(p, alpha=group["weight_decay"])
                p.add_(g, alpha=-group["lr"])
                if self.fuse_zero_grad:
                    p.grad.zero_()
        self._grads_cleared = self.fuse_zero_grad
        return loss


class FusedAdam(_FusedBase):
    """Adam (Kingma & Ba) with torch's default update rule (no amsgrad)."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, fuse_zero_grad: bool = False, bf16_shadow: bool = True,
                 capturable: bool = False) -> None:
        """``capturable=True`` (like ``torch.optim.Adam(capturable=True)``): the step count lives on the device
        (``state[p]["step"]`` is a CUDA tensor shared by the parameters of a group on one device), so ``step()``
        can be captured in a CUDA graph and replayed -- ``Trainer(cuda_graph=True)`` / ``GraphedDataParallel``
        with the reference's Adam loop (01.data_parallel.ipynb:287)."""
        if lr < 0:
            raise ValueError(f"Invalid learning rate: {lr}")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay),
                         fuse_zero_grad, bf16_shadow)
        self.capturable = capturable
        self._step_state: Dict[tuple, torch.Tensor] = {}

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        # the device-side step counters follow the loaded per-parameter counts
        for (gi, dev), ss in self._step_state.items():
            vals = [float(self.state[p]["step"]) for p in self.param_groups[gi]["params"]
                    if p.device == dev and "step" in self.state.get(p, {})]
            ss[0] = max(vals) if vals else 0.0

    def _step_capturable(self, grad_scale: float):
        for gi, group in enumerate(self.param_groups):
            b1, b2 = group["betas"]
            by_dev, slow = self._cuda_groups(group)
            if slow:
                raise RuntimeError("capturable FusedAdam needs contiguous fp32 CUDA parameters and gradients")
            for dev, ps in by_dev.items():
                ss = self._step_state.get((gi, dev))
                if ss is None:
                    if torch.cuda.is_current_stream_capturing():
                        raise RuntimeError("run one eager step() (or call init_state()) before capturing FusedAdam")
                    # resume from a host-side count if this optimizer was stepped / loaded before
                    t0 = max([float(self.state[p]["step"]) for p in ps if "step" in self.state[p]] or [0.0])
                    ss = torch.tensor([t0, 0.0, 0.0], dtype=torch.float32, device=dev)
                    self._step_state[(gi, dev)] = ss
                rows = []
                for p in ps:
                    st = self.state[p]
                    if "exp_avg" not in st:
                        if torch.cuda.is_current_stream_capturing():
                            raise RuntimeError("run one eager step() before capturing FusedAdam")
                        st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["step"] = ss[0]                                     # a view: always the device-side count
                    sh = self._shadow_for(p)
                    rows.append([p.data_ptr(), p.grad.data_ptr(), sh.data_ptr() if sh is not None else 0,
                                 p.numel(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), 0, 0])
                tab = self._tables.setdefault((gi, dev), _DeviceTable(dev))
                tab.update(rows)
                with torch.cuda.device(dev):
                    check(lib().b200dt_multi_adam_capturable(ptr(tab.meta), ptr(tab.chunks), tab.n_chunks, group["lr"],
                                                             b1, b2, group["eps"], group["weight_decay"], ptr(ss),
                                                             grad_scale, int(self.fuse_zero_grad), stream_ptr(dev)),
                          "multi_adam_capturable")
                self._publish_shadows(ps)
        self._grads_cleared = self.fuse_zero_grad

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.capturable:
            self._step_capturable(grad_scale)
            return loss
        if torch.cuda.is_available() and torch.cuda.is_current_stream_capturing():
            raise RuntimeError("this FusedAdam keeps its step count on the host; construct it with "
                               "capturable=True to capture step() in a CUDA graph")
        for gi, group in enumerate(self.param_groups):
            b1, b2 = group["betas"]
            t = None
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if "step" not in st:
                    st["step"] = torch.tensor(0.0)
                st["step"] += 1
                t = int(st["step"].item()) if t is None else t
            if t is None:
                continue
            by_dev, slow = self._cuda_groups(group)
            for dev, ps in by_dev.items():
                rows = []
                for p in ps:
                    st = self.state[p]
                    if "exp_avg" not in st:
                        st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    sh = self._shadow_for(p)
                    rows.append([p.data_ptr(), p.grad.data_ptr(), sh.data_ptr() if sh is not None else 0,
                                 p.numel(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), 0, 0])
                tab = self._tables.setdefault((gi, dev), _DeviceTable(dev))
                tab.update(rows)
                with torch.cuda.device(dev):
                    check(lib().b200dt_multi_adam(ptr(tab.meta), ptr(tab.chunks), tab.n_chunks, group["lr"], b1,
                                                  b2, group["eps"], group["weight_decay"], t, grad_scale,
                                                  int(self.fuse_zero_grad), stream_ptr(dev)), "multi_adam")
                self._publish_shadows(ps)
            for p in slow:
                st = self.state[p]
                if "exp_avg" not in st:
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                g = p.grad * grad_scale if grad_scale != 1.0 else p.grad
                if group["weight_decay"]:
                    g = g.add(p, alpha=group["weight_decay"])
                st["exp_avg"].mul_(b1).add_(g, alpha=1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
                bc1 = 1 - b1 ** t
                bc2 = 1 - b2 ** t
                denom = (st["exp_avg_sq"].sqrt() / (bc2 ** 0.5)).add_(group["eps"])
                p.addcdiv_(st["exp_avg"], denom, value=-group["lr"] / bc1)
                if self.fuse_zero_grad:
                    p.grad.zero_()
        self._grads_cleared = self.fuse_zero_grad
        return loss
