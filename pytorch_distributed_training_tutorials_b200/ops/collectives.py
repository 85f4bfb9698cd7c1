"""Symmetric-memory arena and launch helpers for the fused cross-GPU collectives.

The arena is one peer-mapped staging buffer plus one signal pad per rank.  Two flavours with the same
surface: ``arena.DriverArena`` (the library's own cuMem / cuMulticast bring-up, ``csrc/arena.cu``; the
default, and the one that gives the NVLS kernels their multicast address) and ``SymmetricArena`` on top of
``torch.distributed._symmetric_memory`` (fallback; ``B200DT_ARENA=torch`` selects it).  The kernels that use
the pointers are ours either way (``csrc/allreduce.cu``).  ``TensorTable`` turns a list of tensors (one DDP
bucket) into the device-side pointer / chunk tables that a single ``b200dt_coll_launch`` consumes.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from .. import _lib
from .._lib import check, lib, ptr, u64_array

SIGNAL_WORDS = 8192
MODE = {"oneshot": 0, "twoshot": 1, "nvls": 2, "bcast": 3}
CHUNK_ELEMS = 16384
ONESHOT_MAX_BYTES = 256 * 1024


class CollectiveTimeout(RuntimeError):
    """A peer-memory collective waited longer than B200DT_COLL_TIMEOUT_S for a peer."""


class StatusWords:
    """The collectives' status words: 64 bytes of pinned host memory mapped into the device.  The kernels
    write ``{code, cta << 8 | peer, waited ms}`` there before they trap; the host reads them without
    touching any stream (so it works after the CUDA context has been poisoned by the trap)."""

    _CODES = {1: "a peer never consumed this rank's previous flag (it is stuck or dead)",
              2: "a peer's flag never arrived (it died, hung, or issued its collectives in another order)",
              3: "a peer timed out first and raised the abort flag"}

    def __init__(self) -> None:
        h = lib()._handle
        h.b200dt_status_alloc.argtypes = [C.POINTER(C.POINTER(C.c_int)), C.POINTER(C.c_ulonglong)]
        h.b200dt_status_free.argtypes = [C.POINTER(C.c_int)]
        self._host = C.POINTER(C.c_int)()
        dev = C.c_ulonglong()
        check(h.b200dt_status_alloc(C.byref(self._host), C.byref(dev)), "status_alloc")
        self.dev_ptr = int(dev.value)

    def code(self) -> int:
        return int(self._host[0])

    def check(self, rank: int = -1) -> None:
        c = self.code()
        if c:
            where = int(self._host[1])
            raise CollectiveTimeout(
                f"rank {rank}: peer-memory collective gave up after {int(self._host[2])} ms waiting for rank "
                f"{where & 0xFF} (CTA {where >> 8}): {self._CODES.get(c, 'status ' + str(c))}. The kernel trapped "
                "instead of reducing partial data; this process's CUDA context is no longer usable.")

    def clear(self) -> None:
        self._host[0] = 0


class _SignalRanges:
    """Hands out disjoint ranges of signal words: every stream-level user of an arena (a DDP instance's
    bucket stream, its broadcast stream, a free-standing ``all_reduce_``) gets its own, so launches that may
    overlap on the GPU never share flags.  All ranks must request ranges in the same order."""

    def __init__(self, total: int) -> None:
        self.total = total - 1        # the last word is the abort flag
        self.next = 0

    def take(self, words: int) -> int:
        words = (int(words) + 31) // 32 * 32
        if self.next + words > self.total:
            raise RuntimeError(f"out of signal words ({self.next} + {words} > {self.total}); create a second arena")
        base = self.next
        self.next += words
        return base


def signal_base(arena, key: str, max_blocks: int = 32) -> int:
    """The signal-word range of user ``key`` on ``arena`` (allocated on first use, same order on all ranks)."""
    table = arena.__dict__.setdefault("_sig_users", {})
    if key not in table:
        ranges = arena.__dict__.setdefault("_sig_ranges", _SignalRanges(SIGNAL_WORDS))
        table[key] = ranges.take(max_blocks * arena.world)
    return table[key]


def make_arena(nbytes: int, device: torch.device, group=None, flavour: Optional[str] = None):
    """``flavour``: "driver" (own cuMem / cuMulticast bring-up), "torch" (torch symmetric memory) or None = env
    ``B200DT_ARENA`` / driver first, torch as fallback.  All ranks take the same decision."""
    flavour = flavour or os.environ.get("B200DT_ARENA", "auto")
    group = group if group is not None else dist.group.WORLD
    if flavour in ("auto", "driver"):
        from .arena import DriverArena
        err = None
        arena = None
        try:
            arena = DriverArena(nbytes, device, group)
        except Exception as e:  # noqa: BLE001 - any failure means "use the other flavour", decided collectively
            err = f"{type(e).__name__}: {e}"
        votes = [None] * dist.get_world_size(group)
        dist.all_gather_object(votes, err, group=group)
        bad = [f"rank {r}: {v}" for r, v in enumerate(votes) if v]
        if not bad:
            return arena
        if arena is not None:
            arena.close()
        if flavour == "driver":
            raise RuntimeError("driver-API arena bring-up failed: " + "; ".join(bad))
        print(f"[b200dt] driver-API arena unavailable ({'; '.join(bad)}); using torch symmetric memory",
              file=sys.stderr, flush=True)
    return SymmetricArena(nbytes, device, group)


class SymmetricArena:
    """Per-rank symmetric staging buffer + signal pad, mapped into every peer of the group."""

    flavour = "torch"

    def __init__(self, nbytes: int, device: torch.device, group=None) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        self.device = device
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        if self.world > 8:
            raise RuntimeError("the peer-memory collectives cover one NVSwitch domain (<= 8 GPUs)")
        nbytes = (int(nbytes) + 4095) // 4096 * 4096
        self.nbytes = nbytes
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.sig = symm_mem.empty(SIGNAL_WORDS, dtype=torch.int32, device=device)
        self.buf.zero_()
        self.sig.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group=self.group)
        self.sig_hdl = symm_mem.rendezvous(self.sig, group=self.group)
        self.buf_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.sig_ptrs = [int(p) for p in self.sig_hdl.buffer_ptrs]
        mc = 0
        try:
            if self.hdl.has_multicast_support():
                mc = int(self.hdl.multicast_ptr)
        except Exception:
            mc = 0
        self.multicast_ptr = mc
        self._bufs_c = u64_array(self.buf_ptrs)
        self._sigs_c = u64_array(self.sig_ptrs)
        self.status_words = StatusWords()
        self.multicast_error = None if mc else "torch symmetric memory reports no multicast support"
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group, device_ids=[device.index])   # every pad is zero before first use

    @property
    def has_multicast(self) -> bool:
        return self.multicast_ptr != 0

    @property
    def status(self):
        return self.status_words

    def check_status(self) -> None:
        self.status_words.check(self.rank)

    def close(self) -> None:
        self.hdl = self.sig_hdl = None
        self.buf = self.sig = None


# Measured on 4 x B200 (profiles/allreduce_sweep.md): with fp32 on the wire the two-shot kernel beats the NVLS
# kernel at every size (1 GiB in place: 2.42 vs 3.72 ms), with bf16 on the wire NVLS is the fastest variant
# from 4 MiB up.  "auto" follows the measurements; B200DT_PREFER_NVLS=1/0 overrides it.
NVLS_MIN_BYTES_BF16 = 4 << 20


def pick_mode(nbytes: int, arena, requested: str = "auto", wire_bf16: bool = False) -> str:
    if requested != "auto":
        if requested == "nvls" and not arena.has_multicast:
            raise RuntimeError("allreduce='nvls' requested but the arena has no multicast mapping "
                               f"({getattr(arena, 'multicast_error', None)})")
        return requested
    # one-shot pulls (world-1) x the payload into every rank: only worth it for latency-bound
    # messages, and the break-even shrinks with the number of peers (measured: profiles/)
    if nbytes <= (ONESHOT_MAX_BYTES if arena.world <= 2 else 16 * 1024):
        return "oneshot"
    prefer = os.environ.get("B200DT_PREFER_NVLS", "")
    if arena.has_multicast and arena.world > 2:
        if prefer == "1" or (prefer != "0" and wire_bf16 and nbytes >= NVLS_MIN_BYTES_BF16):
            return "nvls"
    return "twoshot"


class TensorTable:
    """Device-side {ptr, numel} + chunk tables for a fixed list of same-dtype tensors."""

    def __init__(self, numels: Sequence[int], device: torch.device, wire_bf16: bool, tensor_bf16: bool = False,
                 target_blocks: int = 32) -> None:
        self.device = device
        self.numels = [int(n) for n in numels]
        self.tensor_bf16 = tensor_bf16
        self.wire_bf16 = wire_bf16 or tensor_bf16
        # chunk size adapts to the payload: mid-size buckets are spread over all CTAs instead of
        # leaving most of them idle behind a few 64 KB chunks
        total = sum(self.numels)
        chunk = CHUNK_ELEMS
        while chunk > 2048 and total // chunk < target_blocks:
            chunk //= 2
        # ... and large payloads use bigger chunks, so the per-chunk table fetch (two dependent
        # global loads) is amortised over megabytes instead of 64 KB
        while chunk < (1 << 20) and total // chunk > 8 * target_blocks:
            chunk *= 2
        self.chunk_elems = chunk
        rows = []
        off = 0
        for ti, n in enumerate(self.numels):
            s = 0
            while s < n:
                ln = min(chunk, n - s)
                rows.append((ti, s, ln, off))
                off += (ln + 7) // 8 * 8
                s += ln
        self.stage_elems = off
        self.stage_bytes = off * (2 if self.wire_bf16 else 4)
        # element offset of every tensor inside the staging region (its chunks are contiguous)
        self.tensor_offsets = []
        seen = set()
        for ti, s0, ln, o in rows:
            if ti not in seen:
                seen.add(ti)
                self.tensor_offsets.append(o)
        self.n_chunks = len(rows)
        assert lib().b200dt_coll_chunk_bytes() == 24
        raw = torch.empty(max(len(rows), 1) * 3, dtype=torch.int64)
        for i, (ti, s, ln, o) in enumerate(rows):
            raw[3 * i] = (s << 32) | ti                 # little-endian: {int32 tensor, int32 start}
            raw[3 * i + 1] = ln                          # {int32 len, int32 pad}
            raw[3 * i + 2] = o
        self.chunks = raw.to(device)
        self._host = torch.empty(len(self.numels) * 2, dtype=torch.int64)
        self._uploader = _lib.PinnedUploader()
        self.tensors = torch.zeros(len(self.numels) * 2, dtype=torch.int64, device=device)
        self._ptr_key = None

    def grid(self, max_blocks: int) -> int:
        return max(1, min(max_blocks, self.n_chunks))

    def bind(self, tensors: Sequence[torch.Tensor], stream: torch.cuda.Stream) -> None:
        """Point the table at these tensors (skips the upload when the addresses did not change)."""
        key = tuple(t.data_ptr() for t in tensors)
        if key == self._ptr_key:
            return
        want = torch.bfloat16 if self.tensor_bf16 else None
        for i, t in enumerate(tensors):
            assert t.numel() == self.numels[i] and t.is_contiguous()
            if want is not None and t.dtype != want or (want is None and t.element_size() != 4):
                raise TypeError(f"tensor {i} of this table is {t.dtype}; the table was built for "
                                f"{'bf16' if self.tensor_bf16 else '32-bit'} tensors")
            self._host[2 * i] = t.data_ptr()
            self._host[2 * i + 1] = t.numel()
        self._uploader.upload(self.tensors, self._host, stream)
        self._ptr_key = key


def launch(mode: str, arena, table: TensorTable, scale: float, stream: torch.cuda.Stream,
           root: int = 0, sig_base: Optional[int] = None, max_blocks: int = 32, stage_base: int = 0,
           in_place: bool = False) -> None:
    """``stage_base``: byte offset of this table's staging region inside the arena.  ``in_place``: the
    tensors' home IS that region (see ``stage_views``): no copy-in for resident tensors, no copy-out.
    ``sig_base``: first signal word of the caller's own range (``signal_base``); None = the range of
    the stream the launch goes to."""
    grid = table.grid(max_blocks)
    if sig_base is None:
        sig_base = signal_base(arena, f"stream:{stream.cuda_stream}", max(max_blocks, 32))
    if stage_base + table.stage_bytes > arena.nbytes:
        raise RuntimeError(f"bucket of {table.stage_bytes} bytes at offset {stage_base} does not fit the "
                           f"{arena.nbytes}-byte arena")
    if in_place and mode not in ("twoshot", "nvls"):
        raise ValueError("in-place all-reduce needs the two-shot or NVLS kernel")
    with torch.cuda.device(arena.device):
        check(lib().b200dt_coll_launch(MODE[mode], arena._bufs_c, arena._sigs_c, arena.multicast_ptr if mode in ("nvls", "bcast") else 0,
                                       arena.rank, arena.world, root, ptr(table.tensors), ptr(table.chunks),
                                       table.n_chunks, int(stage_base), int(table.tensor_bf16), int(table.wire_bf16),
                                       scale, sig_base, grid, arena.status_words.dev_ptr, int(in_place),
                                       stream.cuda_stream), f"coll_launch[{mode}]")


def stage_views(arena: SymmetricArena, table: TensorTable, shapes: Sequence, stage_base: int = 0) -> List[torch.Tensor]:
    """Tensors that LIVE in this rank's staging region at the slots ``table`` assigns to them: write the
    data there (or let a kernel produce it there) and ``launch(..., in_place=True)`` reduces it without
    the copy-in / copy-out passes -- the peer-memory analogue of NCCL user-buffer registration."""
    dt = torch.bfloat16 if table.wire_bf16 else torch.float32
    esz = 2 if table.wire_bf16 else 4
    if stage_base % 16:
        raise ValueError("stage_base must be a multiple of 16 bytes")
    region = arena.buf[stage_base:stage_base + table.stage_bytes].view(dt)
    out = []
    for off, n, shp in zip(table.tensor_offsets, table.numels, shapes):
        out.append(region[off:off + n].view(shp))
    return out


def all_reduce_(tensors: List[torch.Tensor], arena: SymmetricArena, mode: str = "auto", scale: float = 1.0,
                wire_bf16: bool = False, stream: Optional[torch.cuda.Stream] = None, max_blocks: int = 32) -> None:
    """In-place sum (times ``scale``) of same-dtype CUDA tensors across the arena's group."""
    dev = arena.device
    stream = stream or torch.cuda.current_stream(dev)
    tb16 = tensors[0].dtype == torch.bfloat16
    table = TensorTable([t.numel() for t in tensors], dev, wire_bf16, tb16, target_blocks=max_blocks)
    table.bind(tensors, stream)
    m = pick_mode(table.stage_bytes, arena, mode, table.wire_bf16)
    launch(m, arena, table, scale, stream, max_blocks=max_blocks)


def broadcast_(tensors: List[torch.Tensor], arena: SymmetricArena, root: int = 0,
               stream: Optional[torch.cuda.Stream] = None, max_blocks: int = 32) -> None:
    """In-place broadcast of fp32 / bf16 CUDA tensors from ``root`` through peer stores."""
    dev = arena.device
    stream = stream or torch.cuda.current_stream(dev)
    tb16 = tensors[0].dtype == torch.bfloat16
    table = TensorTable([t.numel() for t in tensors], dev, False, tb16)
    table.bind(tensors, stream)
    launch("bcast", arena, table, 1.0, stream, root=root, max_blocks=max_blocks)
