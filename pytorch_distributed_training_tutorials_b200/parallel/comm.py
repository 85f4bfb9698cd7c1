"""Single-process multi-device communication primitives with autograd support.

These are the building blocks ``DataParallel`` runs every iteration, the counterparts of
``torch.nn.parallel._functions`` / ``comm`` that the reference reaches through
``nn.DataParallel`` (/root/reference/01.data_parallel.ipynb:275,482,486):

* ``Scatter``    split dim 0 into per-device chunks (backward: gather the chunk gradients);
* ``Gather``     concatenate per-device outputs on one device (backward: scatter);
* ``Broadcast``  replicate the parameters onto every device (backward: reduce-add onto the root).

On CUDA the data moves by this library's peer-memory kernels (``csrc/p2p.cu``): parameters are
packed once and pushed to all peers by ONE store kernel, gradients are summed by ONE pull kernel
on the root; devices are ordered with CUDA events.  With CPU "devices" (used by the unit tests on
machines without GPUs) the same code paths degenerate to plain tensor copies.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from .. import _lib


def _dev(d) -> torch.device:
    if isinstance(d, torch.device):
        return d
    if isinstance(d, int):
        return torch.device("cuda", d)
    return torch.device(d)


def _is_cuda(devs: Sequence[torch.device]) -> bool:
    return all(d.type == "cuda" for d in devs)


_peer_enabled = set()


def ensure_peer_access(devs: Sequence[torch.device]) -> None:
    for a in devs:
        for b in devs:
            if a != b and (a.index, b.index) not in _peer_enabled:
                _lib.check(_lib.lib().b200dt_enable_peer_access(a.index, b.index), f"peer access {a}->{b}")
                _peer_enabled.add((a.index, b.index))


def chunk_sizes(n: int, parts: int) -> List[int]:
    """torch.chunk semantics: ceil(n/parts)-sized pieces, possibly fewer than ``parts``."""
    if n == 0:
        return [0]
    step = -(-n // parts)
    sizes = []
    left = n
    while left > 0:
        s = min(step, left)
        sizes.append(s)
        left -= s
    return sizes


def _copy_to(t: torch.Tensor, dev: torch.device) -> torch.Tensor:
    if t.device == dev:
        return t
    return t.to(dev, non_blocking=True)


class Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, devices, dim, x):
        devices = [_dev(d) for d in devices]
        ctx.dim = dim
        ctx.src = x.device
        sizes = chunk_sizes(x.shape[dim], len(devices))
        pieces = torch.split(x, sizes, dim=dim) if x.shape[dim] > 0 else (x,)
        outs = []
        for piece, d in zip(pieces, devices):
            if d == x.device:
                outs.append(piece.contiguous() if d.type == "cuda" else piece)
            else:
                # stream-ordered peer copy: the chunk is ready on d's current stream
                outs.append(_copy_to(piece.contiguous(), d))
        ctx.out_devices = [o.device for o in outs]
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        return None, None, Gather.apply(ctx.src, ctx.dim, *grads)


class Gather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target, dim, *xs):
        target = _dev(target)
        ctx.target = target
        ctx.dim = dim
        ctx.src_devices = [x.device for x in xs]
        ctx.scalar = all(x.dim() == 0 for x in xs)
        if ctx.scalar:
            # scalars become a vector, one entry per replica (torch warns here; so do we, once)
            xs = tuple(x.view(1) for x in xs)
            _warn_scalar()
        ctx.sizes = [x.shape[dim] for x in xs]
        moved = []
        for x in xs:
            if x.device != target and x.is_cuda:
                _wait_current(target, x.device)
            moved.append(_copy_to(x, target))
        return torch.cat(moved, dim=dim)

    @staticmethod
    def backward(ctx, g):
        pieces = torch.split(g, ctx.sizes, dim=ctx.dim)
        outs = []
        for p, d in zip(pieces, ctx.src_devices):
            p = _copy_to(p.contiguous(), d)
            outs.append(p.view(()) if ctx.scalar else p)
        return (None, None) + tuple(outs)


_scalar_warned = False


def _warn_scalar() -> None:
    global _scalar_warned
    if not _scalar_warned:
        import warnings
        warnings.warn("Was asked to gather along dimension 0, but all input tensors were scalars; "
                      "will instead unsqueeze and return a vector.")
        _scalar_warned = True


def _wait_current(waiter: torch.device, producer: torch.device) -> None:
    """Make ``waiter``'s current stream wait for everything queued on ``producer``'s current stream."""
    if waiter.type != "cuda" or producer.type != "cuda" or waiter == producer:
        return
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(producer))
    torch.cuda.current_stream(waiter).wait_event(ev)


class Broadcast(torch.autograd.Function):
    """params (on devices[0]) -> one copy per device; copy 0 is the original tensor itself."""

    @staticmethod
    def forward(ctx, devices, *params):
        devices = [_dev(d) for d in devices]
        ctx.devices = devices
        ctx.n = len(params)
        ctx.root = params[0].device if params else devices[0]
        if not params:
            return ()
        outs: List[torch.Tensor] = list(params)                     # replica 0 aliases the originals
        copies = broadcast_coalesced(params, devices)
        for dev_copies in copies[1:]:
            outs.extend(dev_copies)
        non_diff = [o for o, p in zip(outs[ctx.n:], list(params) * (len(devices) - 1)) if not p.requires_grad]
        ctx.mark_non_differentiable(*non_diff)
        ctx.needs = [p.requires_grad for p in params]
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        n, devices = ctx.n, ctx.devices
        per_dev = [list(grads[i * n:(i + 1) * n]) for i in range(len(devices))]
        reduced = reduce_add_coalesced(per_dev, ctx.root, ctx.needs)
        return (None,) + tuple(reduced)


def broadcast_coalesced(tensors: Sequence[torch.Tensor], devices: Sequence[torch.device]) -> List[List[torch.Tensor]]:
    """Copy every tensor to every device. Returns ``[per-device list]``; entry 0 is the input list.

    CUDA: per dtype the tensors are packed into one flat buffer and pushed to all peers by a single
    store kernel (one read of the source, ``len(devices)-1`` NVLink writes per 16 bytes); the
    replicas' tensors are views into the destination flats, so there is no unpack pass."""
    devices = [_dev(d) for d in devices]
    out: List[List[Optional[torch.Tensor]]] = [list(tensors)] + [[None] * len(tensors) for _ in devices[1:]]
    if len(devices) == 1 or not tensors:
        return out
    root = tensors[0].device
    if not (_is_cuda(devices) and root.type == "cuda"):
        for di, d in enumerate(devices[1:], start=1):
            for ti, t in enumerate(tensors):
                out[di][ti] = t.detach().clone() if d == t.device else t.detach().to(d)
        return out
    ensure_peer_access(devices)
    by_dtype = {}
    for ti, t in enumerate(tensors):
        by_dtype.setdefault(t.dtype, []).append(ti)
    L = _lib.lib()
    for dt, idxs in by_dtype.items():
        group = [tensors[i].detach() for i in idxs]
        esz = group[0].element_size()
        offs, total = [], 0
        for t in group:
            offs.append(total)
            total += t.numel()
        with torch.cuda.device(root):
            flat = torch.cat([t.reshape(-1) for t in group])       # one pack launch
        dst_flats = []
        for d in devices[1:]:
            with torch.cuda.device(d):
                df = torch.empty(total, dtype=dt, device=d)
            _wait_current(root, d)          # d's allocator may still have work queued on that block
            dst_flats.append(df)
        with torch.cuda.device(root):
            _lib.check(L.b200dt_p2p_broadcast(flat.data_ptr(), _lib.u64_array([f.data_ptr() for f in dst_flats]),
                                              len(dst_flats), total * esz, _lib.stream_ptr(root)), "p2p_broadcast")
        for di, (d, df) in enumerate(zip(devices[1:], dst_flats), start=1):
            _wait_current(d, root)          # replicas on d start after the push has been issued & done
            for ti, t, o in zip(idxs, group, offs):
                out[di][ti] = df[o:o + t.numel()].view(t.shape)
    return out


def reduce_add_coalesced(per_device_grads: List[List[Optional[torch.Tensor]]], root: torch.device,
                         needs: Sequence[bool]) -> List[Optional[torch.Tensor]]:
    """Sum the replicas' gradients onto ``root``; one pull kernel on CUDA."""
    n = len(needs)
    ndev = len(per_device_grads)
    result: List[Optional[torch.Tensor]] = [None] * n
    active = [i for i in range(n) if needs[i] and any(per_device_grads[d][i] is not None for d in range(ndev))]
    if not active:
        return result
    devs = []
    for d in range(ndev):
        g = next((x for x in per_device_grads[d] if x is not None), None)
        devs.append(g.device if g is not None else root)
    cuda = root.type == "cuda" and all(x.type == "cuda" for x in devs)
    dense_f32 = all(per_device_grads[d][i] is not None and per_device_grads[d][i].dtype == torch.float32
                    for d in range(ndev) for i in active)
    if not (cuda and dense_f32 and ndev > 1):
        for i in active:
            acc = None
            for d in range(ndev):
                g = per_device_grads[d][i]
                if g is None:
                    continue
                g = _copy_to(g, root)
                acc = g.clone() if acc is None else acc.add_(g)
            result[i] = acc
        return result
    ensure_peer_access(list(dict.fromkeys([root] + devs)))
    flats = []
    for d in range(ndev):
        with torch.cuda.device(devs[d]):
            flats.append(torch.cat([per_device_grads[d][i].reshape(-1) for i in active]))
        _wait_current(root, devs[d])
    total = flats[0].numel()
    with torch.cuda.device(root):
        out = torch.empty(total, dtype=torch.float32, device=root)
        _lib.check(_lib.lib().b200dt_p2p_reduce_f32(out.data_ptr(), _lib.u64_array([f.data_ptr() for f in flats]),
                                                    ndev, total, 1.0, 0, _lib.stream_ptr(root)), "p2p_reduce")
    for d in range(ndev):
        if devs[d] != root:
            flats[d].record_stream(torch.cuda.current_stream(root))
    off = 0
    for i in active:
        g = per_device_grads[0][i]
        result[i] = out[off:off + g.numel()].view(g.shape)
        off += g.numel()
    return result


class GatherSum(torch.autograd.Function):
    """``sum(cat(outputs))`` without materialising the concatenation for the loss: every replica
    pushes its rows into the gathered tensor on the target AND adds its partial sum into the
    target's scalar in the same kernel; backward hands each replica ``grad * ones`` locally, so no
    activation-sized tensor crosses NVLink on the way back."""

    @staticmethod
    def forward(ctx, target, *xs):
        target = _dev(target)
        ctx.shapes = [x.shape for x in xs]
        ctx.devs = [x.device for x in xs]
        ctx.dtypes = [x.dtype for x in xs]
        cuda = target.type == "cuda" and all(x.is_cuda and x.dtype in (torch.float32, torch.bfloat16) for x in xs)
        if not cuda:
            gathered = torch.cat([_copy_to(x, target) for x in xs], dim=0)
            ctx.mark_non_differentiable(gathered)
            return gathered.sum(), gathered
        ensure_peer_access(list(dict.fromkeys([target] + ctx.devs)))
        rows = sum(x.shape[0] for x in xs)
        with torch.cuda.device(target):
            gathered = torch.empty((rows,) + tuple(xs[0].shape[1:]), dtype=xs[0].dtype, device=target)
            total = torch.zeros((), dtype=torch.float32, device=target)
        L = _lib.lib()
        off = 0
        for x in xs:
            xc = x.contiguous()
            _wait_current(x.device, target)       # the zero-fill of `total` precedes the adds
            dst = gathered[off:off + x.shape[0]]
            with torch.cuda.device(x.device):
                _lib.check(L.b200dt_p2p_gather_rows(xc.data_ptr(), dst.data_ptr(), xc.numel() * xc.element_size(),
                                                    total.data_ptr(), int(x.dtype == torch.float32),
                                                    _lib.stream_ptr(x.device)), "p2p_gather_rows")
            _wait_current(target, x.device)
            off += x.shape[0]
        ctx.mark_non_differentiable(gathered)
        return total.to(xs[0].dtype) if xs[0].dtype != torch.float32 else total, gathered

    @staticmethod
    def backward(ctx, g, _unused):
        outs = []
        for shp, d, dt in zip(ctx.shapes, ctx.devs, ctx.dtypes):
            gd = g.to(d, non_blocking=True).to(dt)
            outs.append(gd.expand(shp))
        return (None,) + tuple(outs)
