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

from typing import List, Optional, Sequence

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
        if len(dst_flats) >= 2 and total >= TWO_LEVEL_MIN_ELEMS:
            _two_level_broadcast(flat, dst_flats, root)
        else:
            with torch.cuda.device(root):
                _lib.check(L.b200dt_p2p_broadcast(flat.data_ptr(), _lib.u64_array([f.data_ptr() for f in dst_flats]),
                                                  len(dst_flats), total * esz, _lib.stream_ptr(root)), "p2p_broadcast")
        for di, (d, df) in enumerate(zip(devices[1:], dst_flats), start=1):
            _wait_current(d, root)          # replicas on d start after the push has been issued & done
            for ti, t, o in zip(idxs, group, offs):
                out[di][ti] = df[o:o + t.numel()].view(t.shape)
    return out


# Above this payload the root's NVLink port, not latency, bounds a root-centric broadcast / reduce
# (N-1 copies of everything leave or enter ONE GPU); the two-level forms move (N-1)/N of the payload
# through every port instead.
TWO_LEVEL_MIN_ELEMS = 1 << 18


def _slices(total: int, parts: int, align: int = 64):
    per = -(-total // parts)
    per = (per + align - 1) // align * align
    out = []
    for i in range(parts):
        lo = min(total, i * per)
        out.append((lo, min(total, lo + per)))
    return out


def _two_level_broadcast(flat: torch.Tensor, dst_flats: List[torch.Tensor], root: torch.device) -> None:
    """Scatter + all-gather over NVSwitch: the root sends slice j only to peer j (one payload leaves
    the root instead of N-1), then the peers exchange slices among themselves in N-2 rounds of
    disjoint (source, destination) pairs.  Stream-ordered peer copies; no host synchronisation."""
    m = len(dst_flats)
    sl = _slices(flat.numel(), m)
    with torch.cuda.device(root):
        for j, df in enumerate(dst_flats):
            lo, hi = sl[j]
            if hi > lo:
                df[lo:hi].copy_(flat[lo:hi], non_blocking=True)
    for r in range(1, m):                       # round r: peer i hands its slice to peer (i + r) % m
        for i in range(m):
            j = (i + r) % m
            lo, hi = sl[i]
            if hi > lo:
                with torch.cuda.device(dst_flats[i].device):
                    dst_flats[j][lo:hi].copy_(dst_flats[i][lo:hi], non_blocking=True)


def _two_level_reduce(flats: List[torch.Tensor], devs: List[torch.device], root: torch.device) -> torch.Tensor:
    """Reduce-scatter + gather: device d sums slice d of every replica's flat gradient with the pull
    kernel (its own port carries (N-1)/N of one payload), the root then collects the reduced slices."""
    n = len(flats)
    total = flats[0].numel()
    sl = _slices(total, n)
    L = _lib.lib()
    with torch.cuda.device(root):
        out = torch.empty(total, dtype=torch.float32, device=root)
    ri = next((i for i, d in enumerate(devs) if d == root), None)
    parts = []
    for d in range(n):
        lo, hi = sl[d]
        if hi <= lo:
            parts.append(None)
            continue
        for k in range(n):
            if k != d:
                _wait_current(devs[d], devs[k])          # everybody's flat gradient is complete
        with torch.cuda.device(devs[d]):
            dst = out[lo:hi] if d == ri else torch.empty(hi - lo, dtype=torch.float32, device=devs[d])
            if d == ri:
                _wait_current(devs[d], root)
            _lib.check(L.b200dt_p2p_reduce_f32(dst.data_ptr(), _lib.u64_array([f.data_ptr() + lo * 4 for f in flats]),
                                               n, hi - lo, 1.0, 0, _lib.stream_ptr(devs[d])), "p2p_reduce")
        parts.append(dst)
    with torch.cuda.device(root):
        for d in range(n):
            if parts[d] is not None and d != ri:
                lo, hi = sl[d]
                out[lo:hi].copy_(parts[d], non_blocking=True)
    # the sources must outlive the pulls issued on the other devices' streams
    for d in range(n):
        for k in range(n):
            if k != d and flats[k].is_cuda:
                flats[k].record_stream(torch.cuda.current_stream(devs[d]))
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
    if ndev >= 3 and total >= TWO_LEVEL_MIN_ELEMS and len(set(devs)) == ndev and root in devs:
        out = _two_level_reduce(flats, devs, root)
    else:
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
