"""Device-side timing helpers: CUDA events on the launching stream, max over ranks."""
from __future__ import annotations

import contextlib
import os
import time
from typing import Callable, Optional

import torch


def flush_l2(device=None, nbytes: int = 192 * 1024 * 1024) -> None:
    """Overwrite a buffer larger than the 126 MB L2 so the next kernel starts cold."""
    buf = getattr(flush_l2, "_buf", None)
    if buf is None or buf.device != torch.device(device or torch.cuda.current_device()) or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device or "cuda")
        flush_l2._buf = buf
    buf.zero_()


def time_cuda(fn: Callable[[], None], iters: int = 20, warmup: int = 3, flush: bool = True, device=None) -> float:
    """Median milliseconds per call of ``fn`` measured with CUDA events (one pair per iteration)."""
    for _ in range(max(warmup, 3)):
        fn()
    torch.cuda.synchronize(device)
    times = []
    for _ in range(iters):
        if flush:
            flush_l2(device)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize(device)
        times.append(e0.elapsed_time(e1))
    times.sort()
    return times[len(times) // 2]


def max_over_ranks(value: float, device=None) -> float:
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX + autograd-profiler range so ncu / torch.profiler both see the region."""
    with torch.autograd.profiler.record_function(name):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            if torch.cuda.is_available():
                torch.cuda.nvtx.range_pop()


TRACE = os.environ.get("B200DT_TRACE", "0") not in ("", "0")


def trace_range(name: str):
    """``nvtx_range(name)`` when ``B200DT_TRACE=1`` (forward / backward / allreduce[i] / step show up
    in ncu, nsys and torch.profiler), a free no-op otherwise: the eager hot path issues ~400 kernels
    per step and every microsecond of host time counts."""
    return nvtx_range(name) if TRACE else contextlib.nullcontext()


class StepTimer:
    """Accumulates device time per step; ``summary()`` gives ms/step and samples/s."""

    def __init__(self, device=None) -> None:
        self.device = device
        self.events = []

    @contextlib.contextmanager
    def step(self):
        if torch.cuda.is_available():
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            yield
            e1.record()
            self.events.append((e0, e1))
        else:
            t0 = time.perf_counter()
            yield
            self.events.append(time.perf_counter() - t0)

    def summary(self, samples_per_step: Optional[int] = None) -> dict:
        if torch.cuda.is_available():
            torch.cuda.synchronize(self.device)
            ms = [a.elapsed_time(b) for a, b in self.events]
        else:
            ms = [t * 1e3 for t in self.events]
        if not ms:
            return {}
        mean = sum(ms) / len(ms)
        out = {"steps": len(ms), "ms_per_step": mean}
        if samples_per_step:
            out["samples_per_s"] = samples_per_step / (mean / 1e3)
        return out
