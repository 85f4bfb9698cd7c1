"""Training loop with the reference's call order and console banner.

``Trainer(model, train_dataloader, optimizer, gpu_id)`` moves the model to its device *then* wraps
it in ``DistributedDataParallel`` and, per epoch, probes the batch size, prints
``[GPU: r Epoch: e, Batch size: b | Steps s]``, reseeds the sampler and runs the batches
(/root/reference/ddp_gpus.py:19-53; the torchrun flavour reads ``LOCAL_RANK``,
ddp_gpus_torchrun.py:24).  The banner is byte-for-byte the reference's: its step counts are the
functional oracle (16 steps per epoch on 4 ranks, 64 on one; 02.ddp_toy_example.ipynb:231,268).
"""
from __future__ import annotations

import os
from typing import Callable, Optional

import torch
import torch.nn.functional as F
from torch.utils.data import DataLoader


from .utils.timing import trace_range


def _parse_fault(spec: str):
    """``B200DT_FAULT_INJECT="rank=1,step=3"``: that rank raises at that global step (tests of the
    launchers' failure handling: the first failing child is reported, its siblings are killed)."""
    if not spec:
        return None
    kv = dict(item.split("=") for item in spec.replace(" ", "").split(",") if "=" in item)
    try:
        return int(kv["rank"]), int(kv["step"])
    except (KeyError, ValueError):
        raise ValueError(f"B200DT_FAULT_INJECT must look like 'rank=R,step=S', got {spec!r}")


def _current_rank() -> int:
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank()
    return int(os.environ.get("RANK", "0"))


def default_loss(output: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """``F.cross_entropy`` as the reference calls it (ddp_gpus.py:37); on CUDA the fused kernel."""
    if output.is_cuda and output.dtype in (torch.float32, torch.bfloat16):
        from .ops import functional as Fn
        return Fn.cross_entropy(output, target)
    return F.cross_entropy(output, target)


class Trainer:
    def __init__(self, model: torch.nn.Module, train_dataloader: DataLoader,
                 optimizer: torch.optim.Optimizer, gpu_id: Optional[int] = None,
                 loss_fn: Optional[Callable] = None, wrap_ddp: bool = True, ddp_kwargs: Optional[dict] = None,
                 log_every: int = 0, cuda_graph: bool = False) -> None:
        if gpu_id is None:
            gpu_id = int(os.environ.get("LOCAL_RANK", "0"))
        self.gpu_id = gpu_id
        self.device = torch.device("cuda", gpu_id) if torch.cuda.is_available() else torch.device("cpu")
        self.model = model.to(self.device)
        self.train_dataloader = train_dataloader
        self.optimizer = optimizer
        self.loss_fn = loss_fn or default_loss
        self.log_every = log_every
        self.last_loss: Optional[torch.Tensor] = None
        self.global_step = 0
        self.cuda_graph = cuda_graph and self.device.type == "cuda"
        self._graphed = None
        self._fault = _parse_fault(os.environ.get("B200DT_FAULT_INJECT", ""))
        if wrap_ddp:
            from .parallel.ddp import DistributedDataParallel
            ids = [gpu_id] if self.device.type == "cuda" else None
            self.model = DistributedDataParallel(self.model, device_ids=ids, **(ddp_kwargs or {}))
        if self.cuda_graph and getattr(self.model, "world_size", 1) > 1 and not getattr(self.model, "_use_kernels", True):
            # the NCCL / gloo fallback of the reducer is host-driven (async work handles, waits): it is
            # not captured; only the peer-memory kernels replay from a graph
            import warnings
            warnings.warn("cuda_graph=True needs the fused peer-memory all-reduce; running eager steps instead")
            self.cuda_graph = False

    def _run_batch(self, xs, ys):
        try:
            return self._run_batch_inner(xs, ys)
        except RuntimeError as e:
            self._raise_if_collective_timed_out(e)
            raise

    def _raise_if_collective_timed_out(self, cause: Optional[BaseException] = None) -> None:
        """A peer-memory collective that gives up on a peer traps its kernel; what the host sees next is a
        generic CUDA error.  The status words (pinned host memory) say what really happened."""
        arena = getattr(self.model, "_arena", None)
        if arena is None:
            return
        try:
            arena.check_status()
        except RuntimeError as timeout:
            raise timeout from cause

    def _run_batch_inner(self, xs, ys):
        if self._fault is not None and self._fault == (_current_rank(), self.global_step):
            raise RuntimeError(f"fault injected on rank {self._fault[0]} at step {self._fault[1]} (B200DT_FAULT_INJECT)")
        if self.cuda_graph:
            if self._graphed is None:
                from .utils.graph import GraphedTrainStep
                self._graphed = GraphedTrainStep(self.model, self.loss_fn, self.optimizer, xs, ys)
            if self._graphed.matches(xs, ys):
                self.last_loss = self._graphed(xs, ys)
                self.global_step += 1
                return self.last_loss
        self.optimizer.zero_grad()
        with trace_range("forward"):
            output = self.model(xs)
            loss = self.loss_fn(output, ys)
        with trace_range("backward"):
            loss.backward()
        with trace_range("optimizer.step"):
            self.optimizer.step()
        self.last_loss = loss.detach()
        self.global_step += 1
        return self.last_loss

    def step(self, xs: torch.Tensor, ys: torch.Tensor) -> float:
        """One end-to-end step from host tensors: H2D copy, fwd/bwd/update, loss read back."""
        if self._graphed is not None and self._graphed.matches(xs, ys):
            # copy straight from (pinned) host memory into the graph's static input buffers
            self.last_loss = self._graphed(xs, ys)
            self.global_step += 1
            return float(self.last_loss.item())
        xs = xs.to(self.device, non_blocking=True)
        ys = ys.to(self.device, non_blocking=True)
        try:
            return float(self._run_batch(xs, ys).item())
        except RuntimeError as e:
            self._raise_if_collective_timed_out(e)
            raise

    def run_batches(self, batches) -> list:
        """Train on an iterable of host ``(xs, ys)`` batches, software-pipelined: the upload of batch
        i+1 (``DevicePrefetcher``, copy stream) overlaps the kernels of batch i, and the loss of batch i
        is copied to pinned memory on the compute stream and read on the host while batch i+1 runs.
        Returns every step's loss as a float, in order."""
        from .data import DevicePrefetcher
        losses, pending = [], []
        cuda = self.device.type == "cuda"
        if cuda and getattr(self, "_loss_ring", None) is None:
            self._loss_ring = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        for k, (xs, ys) in enumerate(DevicePrefetcher(batches, self.device)):
            loss = self._run_batch(xs, ys)
            if not cuda:
                losses.append(float(loss))
                continue
            slot = self._loss_ring[k % 2]
            slot.copy_(loss.detach().reshape(1), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            pending.append((slot, ev))
            if len(pending) > 1:
                s, e = pending.pop(0)
                e.synchronize()
                losses.append(float(s[0]))
        for s, e in pending:
            e.synchronize()
            losses.append(float(s[0]))
        return losses

    def _run_epoch(self, epoch: int):
        batch_size = len(next(iter(self.train_dataloader))[0])
        print(f"[GPU: {self.gpu_id} Epoch: {epoch}, Batch size: {batch_size} | Steps {len(self.train_dataloader)}]")
        sampler = getattr(self.train_dataloader, "sampler", None)
        if hasattr(sampler, "set_epoch"):
            sampler.set_epoch(epoch)
        from .data import DevicePrefetcher
        for xs, ys in DevicePrefetcher(self.train_dataloader, self.device):
            self._run_batch(xs, ys)
            if self.log_every and self.global_step % self.log_every == 0:
                from .utils.logging import log_json
                log_json(step=self.global_step, epoch=epoch, loss=float(self.last_loss))

    def train(self, max_epoch: int):
        for epoch in range(max_epoch):
            self._run_epoch(epoch)
