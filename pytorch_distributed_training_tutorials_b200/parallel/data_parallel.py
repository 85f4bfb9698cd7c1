"""``DataParallel``: single-process, multi-GPU data parallelism.

Every forward: scatter the inputs along dim 0, replicate the module onto the devices, run the
replicas in one thread per device, gather the outputs on ``output_device``; backward sums the
replicas' parameter gradients onto the source device.  This is the workflow the reference shows at
/root/reference/01.data_parallel.ipynb:275-277,478-487 (a ``[32,32]`` batch is seen as 4 x
``[8,32]`` inside the replicas and gathers to ``[32,2]``, ibid. 300-473).

B200-first differences from ``torch.nn.DataParallel``: parameters travel as one packed push over
NVLink peer stores and replica parameters are views into the received buffer; gradients come back
through one pull-reduce kernel; worker threads pin their CUDA device *before* the first library
call (the reference's run shows what happens otherwise: the cuBLAS "no current CUDA context"
warning at 01.data_parallel.ipynb:315); ``forward_sum`` fuses the output gather with a sum loss; the
replicas' parameters are re-broadcast only after something wrote the originals (version counters).

Credit: the container recursion of ``scatter`` / ``gather`` below (tensor / namedtuple / tuple / list / dict /
other, in that order, and the "All dicts must have the same number of keys" check) follows PyTorch's
``torch/nn/parallel/scatter_gather.py`` (BSD-3-Clause, (c) the PyTorch contributors) closely on purpose: these
semantics are the API contract ``nn.DataParallel`` users rely on, and the CPU tests compare against it case by case.
"""
from __future__ import annotations

import threading
from typing import Any, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import comm
from .comm import Broadcast, Gather, GatherSum, Scatter, _dev


def _is_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


def scatter(inputs, target_gpus: Sequence, dim: int = 0):
    """Slice tensors into per-device chunks; containers are recursed, other objects are shared."""
    def scatter_map(obj):
        if isinstance(obj, torch.Tensor):
            return list(Scatter.apply(list(target_gpus), dim, obj))
        if _is_namedtuple(obj):
            return [type(obj)(*args) for args in zip(*map(scatter_map, obj))]
        if isinstance(obj, tuple) and len(obj) > 0:
            return list(zip(*map(scatter_map, obj)))
        if isinstance(obj, list) and len(obj) > 0:
            return [list(i) for i in zip(*map(scatter_map, obj))]
        if isinstance(obj, dict) and len(obj) > 0:
            return [type(obj)(i) for i in zip(*map(scatter_map, obj.items()))]
        return [obj for _ in target_gpus]
    return scatter_map(inputs)


def scatter_kwargs(inputs: tuple, kwargs: Optional[dict], target_gpus: Sequence, dim: int = 0):
    """Scatter positional and keyword arguments; both lists are padded to the same length."""
    s_in = scatter(inputs, target_gpus, dim) if inputs else []
    s_kw = scatter(kwargs, target_gpus, dim) if kwargs else []
    if len(s_in) < len(s_kw):
        s_in.extend(() for _ in range(len(s_kw) - len(s_in)))
    elif len(s_kw) < len(s_in):
        s_kw.extend({} for _ in range(len(s_in) - len(s_kw)))
    return tuple(s_in), tuple(s_kw)


def gather(outputs: Sequence, target_device, dim: int = 0):
    """Inverse of ``scatter`` for module outputs (tensors, None, dicts, tuples, namedtuples)."""
    def gather_map(outs):
        out = outs[0]
        if isinstance(out, torch.Tensor):
            return Gather.apply(target_device, dim, *outs)
        if out is None:
            return None
        if isinstance(out, dict):
            if not all(len(out) == len(d) for d in outs):
                raise ValueError("All dicts must have the same number of keys")
            return type(out)((k, gather_map([d[k] for d in outs])) for k in out)
        if _is_namedtuple(out):
            return type(out)._make(map(gather_map, zip(*outs)))
        return type(out)(map(gather_map, zip(*outs)))
    return gather_map(list(outputs))


def replicate(module: nn.Module, devices: Sequence, detach: bool = False,
              cache: Optional[dict] = None) -> List[nn.Module]:
    """Functional copies of ``module`` on each device. Replica 0 shares storage with the original
    (in-place buffer updates such as BatchNorm statistics survive); the other replicas hold
    non-leaf parameter tensors produced by ``Broadcast`` so gradients flow back to the originals."""
    devices = [_dev(d) for d in devices]
    n = len(devices)
    params = list(module.parameters())
    p_index = {p: i for i, p in enumerate(params)}
    pcache = cache.setdefault("params", {}) if cache is not None else None
    bcache = cache.setdefault("buffers", {}) if cache is not None else None
    if params:
        if detach:
            per_dev = comm.broadcast_coalesced(params, devices, pcache)
            per_dev = [[t.detach() for t in lst] for lst in per_dev]
        else:
            flat = Broadcast.apply((devices, pcache) if pcache is not None else devices, *params)
            per_dev = [list(flat[i * len(params):(i + 1) * len(params)]) for i in range(n)]
    else:
        per_dev = [[] for _ in range(n)]
    buffers = list(module.buffers())
    b_index = {b: i for i, b in enumerate(buffers)}
    if buffers:
        with torch.no_grad():
            b_per_dev = comm.broadcast_coalesced(buffers, devices, bcache)
        b_per_dev[0] = list(buffers)
    else:
        b_per_dev = [[] for _ in range(n)]
    modules = list(module.modules())
    m_index = {m: i for i, m in enumerate(modules)}
    copies: List[List[nn.Module]] = [[] for _ in range(n)]
    for m in modules:
        for j in range(n):
            copies[j].append(m._replicate_for_data_parallel())
    for i, m in enumerate(modules):
        for key, child in m._modules.items():
            for j in range(n):
                r = copies[j][i]
                r._modules[key] = None if child is None else copies[j][m_index[child]]
        for key, p in m._parameters.items():
            for j in range(n):
                r = copies[j][i]
                if p is None:
                    r._parameters[key] = None
                else:
                    # plain attribute (not nn.Parameter): replicas are functional views of the owner
                    setattr(r, key, per_dev[j][p_index[p]])
        for key, b in m._buffers.items():
            for j in range(n):
                r = copies[j][i]
                if b is None:
                    r._buffers[key] = None
                else:
                    setattr(r, key, b_per_dev[j][b_index[b]])
    return [copies[j][0] for j in range(n)]


def parallel_apply(modules: Sequence[nn.Module], inputs: Sequence, kwargs_tup: Optional[Sequence[dict]] = None,
                   devices: Optional[Sequence] = None, threaded: bool = True) -> List[Any]:
    """Run ``modules[i](*inputs[i], **kwargs[i])`` concurrently, one thread per replica.

    ``threaded=False`` issues the replicas one after the other from the calling thread instead.  On
    CUDA devices that is still concurrent execution -- every launch is asynchronous and each replica
    has its own device -- without thread start/join and GIL hand-offs, which cost more than the five
    launches of a small model take to issue (measured: a 2-GPU MLP step drops from 3.6 to ~2 ms)."""
    assert len(modules) == len(inputs)
    kwargs_tup = kwargs_tup if kwargs_tup is not None else tuple({} for _ in modules)
    devices = [_dev(d) for d in devices] if devices is not None else [None] * len(modules)
    results: dict = {}
    lock = threading.Lock()
    grad_enabled = torch.is_grad_enabled()
    streams = [torch.cuda.current_stream(d) if d is not None and d.type == "cuda" else None for d in devices]

    def worker(i: int, module: nn.Module, inp, kw, dev, stream) -> None:
        try:
            torch.set_grad_enabled(grad_enabled)
            if not isinstance(inp, (list, tuple)):
                inp = (inp,)
            if dev is not None and dev.type == "cuda":
                # make the device current for this thread before ANY CUDA library call
                with torch.cuda.device(dev), torch.cuda.stream(stream):
                    out = module(*inp, **kw)
            else:
                out = module(*inp, **kw)
            with lock:
                results[i] = out
        except Exception as e:
            with lock:
                results[i] = _Failure(e, i, dev)

    if len(modules) > 1 and not threaded:
        for i, (m, inp, kw, d, s) in enumerate(zip(modules, inputs, kwargs_tup, devices, streams)):
            worker(i, m, inp, kw, d, s)
            if isinstance(results[i], _Failure):
                break
    elif len(modules) > 1:
        threads = [threading.Thread(target=worker, args=(i, m, inp, kw, d, s))
                   for i, (m, inp, kw, d, s) in enumerate(zip(modules, inputs, kwargs_tup, devices, streams))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    else:
        worker(0, modules[0], inputs[0], kwargs_tup[0], devices[0], streams[0])
    outs = []
    for i in range(len(modules)):
        r = results.get(i)
        if isinstance(r, _Failure):
            r.reraise()
        outs.append(r)
    return outs


class _Failure:
    def __init__(self, exc: Exception, index: int, dev) -> None:
        import traceback
        self.exc = exc
        self.msg = f"Caught {type(exc).__name__} in replica {index} on device {dev}.\n" + \
            "".join(traceback.format_exception(type(exc), exc, exc.__traceback__))

    def reraise(self):
        try:
            raise type(self.exc)(self.msg)
        except TypeError:
            raise RuntimeError(self.msg) from self.exc


class DataParallel(nn.Module):
    """``threaded``: one Python thread per replica (``torch.nn.DataParallel``'s way).  Default ``None``:
    threads on CPU placeholder devices, in-order asynchronous issue from the calling thread on CUDA."""

    def __init__(self, module: nn.Module, device_ids: Optional[Sequence] = None, output_device=None, dim: int = 0,
                 threaded: Optional[bool] = None) -> None:
        super().__init__()
        self.threaded = threaded
        # replicas of the parameters / buffers survive between forwards until a source tensor is written
        # (version counters): no re-broadcast for evaluation loops or several forwards per optimizer step
        self._replica_cache: dict = {}
        if device_ids is None:
            device_ids = list(range(torch.cuda.device_count())) if torch.cuda.is_available() else []
        self.module = module
        self.dim = dim
        self.device_ids = [_dev(d) for d in device_ids]
        if output_device is None and self.device_ids:
            output_device = self.device_ids[0]
        self.output_device = _dev(output_device) if output_device is not None else None
        self.src_device = self.device_ids[0] if self.device_ids else None
        if len(self.device_ids) == 1 and self.src_device.type == "cuda":
            self.module.to(self.src_device)

    def _check_source(self) -> None:
        for t in list(self.module.parameters()) + list(self.module.buffers()):
            if t.device != self.src_device:
                raise RuntimeError("module must have its parameters and buffers on device "
                                   f"{self.src_device} (device_ids[0]) but found one of them on device: {t.device}")

    def _run_replicas(self, inputs, kwargs):
        self._check_source()
        s_in, s_kw = scatter_kwargs(inputs, kwargs, self.device_ids, self.dim)
        if not s_in and not s_kw:
            s_in, s_kw = ((),), ({},)
        if len(self.device_ids) == 1:
            return [self.module(*s_in[0], **s_kw[0])]
        used = self.device_ids[:len(s_in)]
        replicas = replicate(self.module, used, detach=not torch.is_grad_enabled(), cache=self._replica_cache)
        threaded = self.threaded if self.threaded is not None else any(d.type != "cuda" for d in used)
        return parallel_apply(replicas, s_in, s_kw, used, threaded=threaded)

    def forward(self, *inputs, **kwargs):
        with torch.autograd.profiler.record_function("b200dt.DataParallel.forward"):
            if not self.device_ids:
                return self.module(*inputs, **kwargs)
            outs = self._run_replicas(inputs, kwargs)
            if len(outs) == 1 and len(self.device_ids) == 1:
                return outs[0]
            return gather(outs, self.output_device, self.dim)

    def forward_sum(self, *inputs, **kwargs) -> Tuple[torch.Tensor, torch.Tensor]:
        """``(output.sum(), output)`` with the gather fused into the loss reduction (the reference's
        loss is exactly ``output.sum()``, 01.data_parallel.ipynb:485).  ``output`` is detached."""
        if not self.device_ids:
            out = self.module(*inputs, **kwargs)
            return out.sum(), out.detach()
        outs = self._run_replicas(inputs, kwargs)
        return GatherSum.apply(self.output_device, *outs)
