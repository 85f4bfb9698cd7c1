"""Automatic layer placement by byte budget (the capability behind ``device_map="auto"``).

The reference loads LLaMA-7B with ``device_map="auto"`` and lists where every parameter landed
(/root/reference/03.model_parallel.ipynb:52-56,114-404): whole decoder layers stay together, are
assigned to the GPUs in order, the embedding goes with the first layers and the final norm + head
with the last ones.  ``auto_place`` reproduces that policy for any model that exposes its blocks in
execution order; ``dispatch`` moves the blocks and installs hooks that move activations.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from .comm import _dev
from .model_parallel import hop


def module_bytes(m: nn.Module) -> int:
    seen, total = set(), 0
    for t in list(m.parameters()) + list(m.buffers()):
        if id(t) not in seen:
            seen.add(id(t))
            total += t.numel() * t.element_size()
    return total


def execution_blocks(model: nn.Module) -> List[Tuple[str, nn.Module]]:
    """Top-level children in order; ``ModuleList`` / ``Sequential`` children are expanded one
    level so that each repeated layer is its own placement unit."""
    out: List[Tuple[str, nn.Module]] = []
    for name, child in model.named_children():
        if isinstance(child, (nn.ModuleList, nn.Sequential)) and len(child) > 1:
            out.extend((f"{name}.{i}", sub) for i, sub in enumerate(child))
        else:
            out.append((name, child))
    return out


def auto_place(model: nn.Module, devices: Sequence, max_memory: Optional[Dict] = None,
               blocks: Optional[List[Tuple[str, nn.Module]]] = None) -> Dict[str, torch.device]:
    """Assign consecutive blocks to devices. Each device takes blocks while its budget allows;
    the last device takes whatever remains. Default budgets: an equal share of the total bytes."""
    devices = [_dev(d) for d in devices]
    blocks = blocks if blocks is not None else execution_blocks(model)
    sizes = [module_bytes(m) for _, m in blocks]
    total = sum(sizes)
    if max_memory is None:
        share = -(-total // len(devices))
        budgets = [share] * len(devices)
    else:
        budgets = [int(max_memory.get(d, max_memory.get(str(d), max_memory.get(d.index, 0)))) for d in devices]
    placement: Dict[str, torch.device] = {}
    di, used = 0, 0
    for (name, _), sz in zip(blocks, sizes):
        while di < len(devices) - 1 and used + sz > budgets[di] and used > 0:
            di, used = di + 1, 0
        placement[name] = devices[di]
        used += sz
    return placement


def dispatch(model: nn.Module, placement: Dict[str, torch.device]) -> nn.Module:
    """Move every placed block to its device and make it pull its inputs over on entry."""
    mods = dict(model.named_modules())
    for name, dev in placement.items():
        block = mods[name]
        block.to(dev)

        def pre_hook(module, args, kwargs, dev=dev):
            def mv(o):
                if isinstance(o, torch.Tensor):
                    return hop(o, dev) if o.is_floating_point() and o.requires_grad else o.to(dev)
                if isinstance(o, (list, tuple)):
                    return type(o)(mv(v) for v in o)
                if isinstance(o, dict):
                    return {k: mv(v) for k, v in o.items()}
                return o
            return mv(args), mv(kwargs)

        block.register_forward_pre_hook(pre_hook, with_kwargs=True)
    model._b200_placement = placement
    return model


def placement_table(model: nn.Module) -> List[Tuple[int, str, torch.dtype, torch.device]]:
    """``(index, name, dtype, device)`` per parameter: the listing the reference prints
    (03.model_parallel.ipynb:409-410)."""
    return [(i, n, p.dtype, p.device) for i, (n, p) in enumerate(model.named_parameters())]


# --------------------------------------------------------------------------- placement by measured time
def profile_units(units: Sequence[Callable], x: torch.Tensor, reps: int = 3,
                  module: Optional[nn.Module] = None, graphed: bool = True) -> List[float]:
    """Device time (ms) of forward + backward of every unit of a chain on ``x``'s device: unit ``k`` is fed the
    output of unit ``k-1``.  Each unit is captured into a CUDA graph and the best of ``reps`` replays is kept, so
    the numbers are what a graph-replayed pipeline will see even where eager launches are host-bound (small
    micro-batches); ``graphed=False`` (or a unit that cannot be captured) times the eager launches instead.
    Buffers of ``module`` (running statistics) are restored and its gradients dropped afterwards.  Byte budgets
    place a LLaMA (``auto_place``); a pipeline wants equal TIME per stage, and for a conv net the two differ a
    lot: ResNet-50's first half holds 6 % of the parameters and takes 70 % of the step."""
    saved = [b.detach().clone() for b in module.buffers()] if module is not None else []
    times: List[float] = []
    cur = x
    if x.device.type != "cuda":
        # CPU placeholders (tests, debugging): wall-clock time of the eager forward + backward
        import time
        for f in units:
            xin = cur.detach().requires_grad_(cur.is_floating_point())
            best, y = float("inf"), None
            for _ in range(max(1, reps)):
                xin.grad = None
                t0 = time.perf_counter()
                y = f(xin)
                y.backward(torch.ones_like(y))
                best = min(best, (time.perf_counter() - t0) * 1e3)
            times.append(best)
            cur = y.detach()
        if module is not None:
            with torch.no_grad():
                for b, s in zip(module.buffers(), saved):
                    b.copy_(s)
            for p in module.parameters():
                p.grad = None
        return times
    with torch.cuda.device(x.device):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        side = torch.cuda.Stream()
        for f in units:
            xin = cur.detach().requires_grad_(cur.is_floating_point())
            y = f(xin)                                   # eager warm-up: lazy initialisation, output shape
            g = torch.ones_like(y)
            y.backward(g)
            torch.cuda.synchronize()
            run = None
            if graphed:
                try:
                    xin.grad = None
                    side.wait_stream(torch.cuda.current_stream())
                    gph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gph, stream=side):
                        yg = f(xin)
                        yg.backward(g)
                    run = gph.replay
                except Exception:                        # noqa: BLE001 -- not capturable: time it eagerly
                    torch.cuda.synchronize()
                    run = None
            if run is None:
                def run(f=f, xin=xin, g=g):
                    xin.grad = None
                    f(xin).backward(g)
            best = float("inf")
            for _ in range(reps + 1):
                torch.cuda.synchronize()
                e0.record()
                run()
                e1.record()
                e1.synchronize()
                if _ > 0:
                    best = min(best, e0.elapsed_time(e1))
            times.append(best)
            cur = y.detach()
    if module is not None:
        with torch.no_grad():
            for b, s in zip(module.buffers(), saved):
                b.copy_(s)
        for p in module.parameters():
            p.grad = None
    return times


def balanced_cuts(times: Sequence[float], n_stages: int) -> List[int]:
    """Indices of the last unit of every stage but the final one, minimising the slowest stage's total
    (contiguous partition of ``times`` into ``n_stages`` non-empty runs; exact, by dynamic programming)."""
    n = len(times)
    if not 1 <= n_stages <= n:
        raise ValueError(f"cannot cut {n} units into {n_stages} stages")
    pre = [0.0]
    for t in times:
        pre.append(pre[-1] + float(t))
    INF = float("inf")
    # best[s][i]: minimal bottleneck for the first i units in s stages
    best = [[INF] * (n + 1) for _ in range(n_stages + 1)]
    arg = [[0] * (n + 1) for _ in range(n_stages + 1)]
    best[0][0] = 0.0
    for s in range(1, n_stages + 1):
        for i in range(s, n + 1):
            for j in range(s - 1, i):
                v = max(best[s - 1][j], pre[i] - pre[j])
                if v < best[s][i]:
                    best[s][i], arg[s][i] = v, j
    cuts, i = [], n
    for s in range(n_stages, 1, -1):
        i = arg[s][i]
        cuts.append(i - 1)
    return cuts[::-1]
