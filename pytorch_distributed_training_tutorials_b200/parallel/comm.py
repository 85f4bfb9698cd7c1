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


def _multi_copy(srcs: Sequence[torch.Tensor], dsts: Sequence[torch.Tensor], on: torch.device) -> None:
    """``dsts[k].copy_(srcs[k])`` for up to 8 contiguous same-size pairs as ONE kernel on ``on``'s current stream
    (peer loads / stores over NVLink; ``csrc/p2p.cu`` ``p2p_multi_copy_kernel``)."""
    import ctypes
    n = len(srcs)
    nbytes = (ctypes.c_longlong * n)(*[t.numel() * t.element_size() for t in srcs])
    with torch.cuda.device(on):
        _lib.check(_lib.lib().b200dt_p2p_multi_copy(_lib.u64_array([t.data_ptr() for t in srcs]),
                                                    _lib.u64_array([t.data_ptr() for t in dsts]), nbytes, n,
                                                    _lib.stream_ptr(on)), "p2p_multi_copy")


def _kernel_scatter_ok(x: torch.Tensor, devices: Sequence[torch.device], dim: int) -> bool:
    return (x.is_cuda and dim == 0 and x.dim() >= 1 and x.is_contiguous() and x.shape[0] > 0 and len(devices) <= 8
            and _is_cuda(devices))


class Scatter(torch.autograd.Function):
    """Split dim 0 into per-device chunks.  CUDA: the chunks of every peer leave the source device in ONE
    store kernel (K5; torch issues one cudaMemcpyPeerAsync per chunk, torch:nn/parallel/comm.py:209)."""

    @staticmethod
    def forward(ctx, devices, dim, x):
        devices = [_dev(d) for d in devices]
        ctx.dim = dim
        ctx.src = x.device
        sizes = chunk_sizes(x.shape[dim], len(devices))
        pieces = torch.split(x, sizes, dim=dim) if x.shape[dim] > 0 else (x,)
        outs = []
        if _kernel_scatter_ok(x, devices, dim) and len(pieces) > 1:
            ensure_peer_access(list(dict.fromkeys([x.device] + devices[:len(pieces)])))
            srcs, dsts = [], []
            for piece, d in zip(pieces, devices):
                if d == x.device:
                    outs.append(piece)                       # a view: nothing moves
                    continue
                with torch.cuda.device(d):
                    o = torch.empty(piece.shape, dtype=x.dtype, device=d)
                _wait_current(x.device, d)                   # d's allocator may still have work queued on that block
                outs.append(o)
                srcs.append(piece); dsts.append(o)
            if srcs:
                _multi_copy(srcs, dsts, x.device)
                for o in dsts:
                    _wait_current(o.device, x.device)        # the chunk is ready on d's current stream
                    o.record_stream(torch.cuda.current_stream(x.device))
        else:
            for piece, d in zip(pieces, devices):
                if d == x.device:
                    outs.append(piece.contiguous() if d.type == "cuda" else piece)
                else:
                    # host input or a non-leading dimension: stream-ordered copy per chunk
                    outs.append(_copy_to(piece.contiguous(), d))
        ctx.out_devices = [o.device for o in outs]
        return tuple(outs)

    @staticmethod
    def backward(ctx, *grads):
        return None, None, Gather.apply(ctx.src, ctx.dim, *grads)


class Gather(torch.autograd.Function):
    """Concatenate per-device outputs on ``target``.  CUDA: the target PULLS every peer's rows straight into its
    slice of the result in ONE kernel (K6; no staging copies, no ``cat``)."""

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
        fast = (target.type == "cuda" and dim == 0 and 1 < len(xs) <= 8 and all(x.is_cuda and x.dim() >= 1 for x in xs)
                and all(x.dtype == xs[0].dtype and x.shape[1:] == xs[0].shape[1:] for x in xs))
        if fast:
            ensure_peer_access(list(dict.fromkeys([target] + [x.device for x in xs])))
            with torch.cuda.device(target):
                out = torch.empty((sum(ctx.sizes),) + tuple(xs[0].shape[1:]), dtype=xs[0].dtype, device=target)
            srcs, dsts, off = [], [], 0
            for x in xs:
                n = x.shape[0]
                if n:
                    xc = x.contiguous()
                    if xc.device != target:
                        _wait_current(target, xc.device)
                        xc.record_stream(torch.cuda.current_stream(target))
                    srcs.append(xc); dsts.append(out[off:off + n])
                off += n
            if srcs:
                _multi_copy(srcs, dsts, target)
            return out
        moved = []
        for x in xs:
            if x.device != target and x.is_cuda:
                _wait_current(target, x.device)
            moved.append(_copy_to(x, target))
        return torch.cat(moved, dim=dim)

    @staticmethod
    def backward(ctx, g):
        devs = ctx.src_devices
        if g.is_cuda and ctx.dim == 0 and g.is_contiguous() and len(devs) > 1 and _is_cuda(devs) and len(devs) <= 8:
            # scatter of the gradient: one store kernel on the target for all peers
            pieces = torch.split(g, ctx.sizes, dim=0)
            ensure_peer_access(list(dict.fromkeys([g.device] + devs)))
            outs, srcs, dsts = [], [], []
            for p, d in zip(pieces, devs):
                if d == g.device:
                    outs.append(p)
                    continue
                with torch.cuda.device(d):
                    o = torch.empty(p.shape, dtype=g.dtype, device=d)
                _wait_current(g.device, d)
                outs.append(o); srcs.append(p); dsts.append(o)
            if srcs:
                _multi_copy(srcs, dsts, g.device)
                for o in dsts:
                    _wait_current(o.device, g.device)
                    o.record_stream(torch.cuda.current_stream(g.device))
            return (None, None) + tuple(o.view(()) if ctx.scalar else o for o in outs)
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


_events = {}


def _wait_current(waiter: torch.device, producer: torch.device) -> None:
    """Make ``waiter``'s current stream wait for everything queued on ``producer``'s current stream.
    One event per (thread, producer, waiter) pair is re-recorded: a wait captures the event's state when it is
    issued, so re-use is safe and saves an event creation per call (~10 of them per DataParallel step)."""
    if waiter.type != "cuda" or producer.type != "cuda" or waiter == producer:
        return
    import threading
    key = (threading.get_ident(), producer.index, waiter.index)
    ev = _events.get(key)
    if ev is None or torch.cuda.is_current_stream_capturing():
        ev = torch.cuda.Event()
        if not torch.cuda.is_current_stream_capturing():
            _events[key] = ev
    ev.record(torch.cuda.current_stream(producer))
    torch.cuda.current_stream(waiter).wait_event(ev)


class Broadcast(torch.autograd.Function):
    """params (on devices[0]) -> one copy per device; copy 0 is the original tensor itself."""

    @staticmethod
    def forward(ctx, devices, *params):
        cache = None
        if isinstance(devices, tuple) and len(devices) == 2 and isinstance(devices[1], dict):
            devices, cache = devices                                  # (devices, replica cache of the caller)
        devices = [_dev(d) for d in devices]
        ctx.devices = devices
        ctx.n = len(params)
        ctx.root = params[0].device if params else devices[0]
        if not params:
            return ()
        outs: List[torch.Tensor] = list(params)                     # replica 0 aliases the originals
        copies = broadcast_coalesced(params, devices, cache)
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


def _version_key(tensors: Sequence[torch.Tensor], devices: Sequence[torch.device]):
    return (tuple(devices), tuple((t.data_ptr(), t._version, t.dtype, tuple(t.shape)) for t in tensors))


def broadcast_coalesced(tensors: Sequence[torch.Tensor], devices: Sequence[torch.device],
                        cache: Optional[dict] = None) -> List[List[torch.Tensor]]:
    """Copy every tensor to every device. Returns ``[per-device list]``; entry 0 is the input list.

    CUDA: per dtype the tensors are packed into one flat buffer and pushed to all peers by a single
    store kernel (one read of the source, ``len(devices)-1`` NVLink writes per 16 bytes); the
    replicas' tensors are views into the destination flats, so there is no unpack pass.

    ``cache`` (a dict owned by the caller, e.g. one per ``DataParallel``): the replicas of the last call are
    kept together with the (address, version counter) of every source tensor; while nothing was written since
    -- no ``optimizer.step()``, no in-place buffer update -- the call moves no data and launches nothing (K4:
    torch's ``replicate`` re-broadcasts on every forward, torch:nn/parallel/replicate.py:108)."""
    devices = [_dev(d) for d in devices]
    out: List[List[Optional[torch.Tensor]]] = [list(tensors)] + [[None] * len(tensors) for _ in devices[1:]]
    if len(devices) == 1 or not tensors:
        return out
    if cache is not None:
        key = _version_key(tensors, devices)
        if cache.get("key") == key:
            cache["hits"] = cache.get("hits", 0) + 1
            return [list(tensors)] + [[c.detach() for c in dev_copies] for dev_copies in cache["copies"][1:]]
        out = broadcast_coalesced(tensors, devices, None)
        cache["key"] = key
        cache["copies"] = [None] + [[c.detach() for c in dev_copies] for dev_copies in out[1:]]
        cache["misses"] = cache.get("misses", 0) + 1
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
        two_level = len(dst_flats) >= 2 and total >= TWO_LEVEL_MIN_ELEMS
        if two_level:
            _two_level_broadcast(flat, dst_flats, root)          # (orders the replicas' streams itself)
        else:
            with torch.cuda.device(root):
                _lib.check(L.b200dt_p2p_broadcast(flat.data_ptr(), _lib.u64_array([f.data_ptr() for f in dst_flats]),
                                                  len(dst_flats), total * esz, _lib.stream_ptr(root)), "p2p_broadcast")
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(root))
        for di, (d, df) in enumerate(zip(devices[1:], dst_flats), start=1):
            if not two_level:
                torch.cuda.current_stream(d).wait_event(done)    # replicas on d start after the push is done
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
    """Scatter + all-gather over NVSwitch in N launches: ONE kernel on the root hands slice j to peer j (one
    payload leaves the root instead of N-1), then ONE kernel per peer pushes its slice to all the other peers
    (N-2 stores per 16 bytes read).  Ordered with events; no host synchronisation."""
    m = len(dst_flats)
    sl = _slices(flat.numel(), m)
    esz = flat.element_size()
    L = _lib.lib()
    srcs = [flat[lo:hi] for (lo, hi) in sl if hi > lo]
    dsts = [df[lo:hi] for df, (lo, hi) in zip(dst_flats, sl) if hi > lo]
    _multi_copy(srcs, dsts, root)
    # one event per producer, any number of waiters.  The peers' destination blocks need no wait of their own:
    # the root waited for every peer's allocator state before the hand-out, and each push waits for the root.
    handed = torch.cuda.Event()
    handed.record(torch.cuda.current_stream(root))
    pushed = {}
    for i, df in enumerate(dst_flats):
        lo, hi = sl[i]
        if hi <= lo:
            continue
        d = df.device
        torch.cuda.current_stream(d).wait_event(handed)          # slice i has arrived on peer i
        others = [o for j, o in enumerate(dst_flats) if j != i]
        with torch.cuda.device(d):
            _lib.check(L.b200dt_p2p_broadcast(df.data_ptr() + lo * esz,
                                              _lib.u64_array([o.data_ptr() + lo * esz for o in others]), len(others),
                                              (hi - lo) * esz, _lib.stream_ptr(d)), "p2p_broadcast")
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(d))
        pushed[i] = ev
    for j, df in enumerate(dst_flats):          # replica j runs after every slice has been pushed into it
        sj = torch.cuda.current_stream(df.device)
        for i, ev in pushed.items():
            if i != j:
                sj.wait_event(ev)
        if j not in pushed:
            sj.wait_event(handed)


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
