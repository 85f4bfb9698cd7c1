"""Symmetric arena brought up on the CUDA driver API by this library itself (``csrc/arena.cu``).

One ``cuMemCreate`` allocation per rank holds the staging buffer and the signal pad; it is exported as a
POSIX file descriptor, handed to every peer process over an ``AF_UNIX`` socket (``SCM_RIGHTS``), mapped
there, and bound to one NVLS multicast object (``cuMulticastCreate`` / ``AddDevice`` / ``BindMem``) so the
``multimem.ld_reduce`` / ``multimem.st`` kernels of ``csrc/allreduce.cu`` have an address to use.  This is the
role ProcessGroupNCCL's buffer registration plays behind /root/reference/ddp_gpus.py:16; torch's own
``_symmetric_memory`` (the other arena flavour, ``collectives.SymmetricArena``) reports no multicast support on
the B200 boxes this was developed on, although the driver offers it (``profiles/nvls_probe.md``).

Control-plane only: rendezvous uses ``torch.distributed`` object collectives of the group.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import uuid
from typing import List

import torch
import torch.distributed as dist

from .._lib import check, lib, u64_array

SIGNAL_WORDS = 8192


class _Foreign:
    """Device memory owned by the arena, exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner) -> None:
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}
        self._owner = owner


def _barrier(group, device=None) -> None:
    if device is not None and device.type == "cuda" and dist.get_backend(group) == "nccl":
        dist.barrier(group=group, device_ids=[device.index])
    else:
        dist.barrier(group=group)


def _exchange_fds(group, rank: int, world: int, my_fd: int, tag: str, only_from: int = -1, device=None) -> List[int]:
    """Every rank publishes one fd; returns the fds of all ranks as received into THIS process (own slot:
    ``my_fd``).  ``only_from >= 0``: only that rank publishes (everyone else receives its fd)."""
    name = f"\0b200dt-{tag}-{rank}"
    srv = None
    serving = only_from < 0 or only_from == rank
    if serving:
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(name)
        srv.listen(world)
    _barrier(group, device)                        # every server socket is listening
    out = [-1] * world
    out[rank] = my_fd
    sources = range(world) if only_from < 0 else [only_from]
    # connect to every publisher first (connections queue in the backlog), then serve our own clients
    conns = {}
    for r in sources:
        if r == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(f"\0b200dt-{tag}-{r}")
        conns[r] = c
    if serving:
        for _ in range(world - 1):
            conn, _addr = srv.accept()
            socket.send_fds(conn, [b"f"], [my_fd])
            conn.close()
    for r, c in conns.items():
        _msg, fds, _flags, _addr = socket.recv_fds(c, 16, 1)
        if not fds:
            raise RuntimeError(f"rank {rank}: no file descriptor received from rank {r}")
        out[r] = fds[0]
        c.close()
    if srv is not None:
        srv.close()
    return out


class DriverArena:
    """Same surface as ``collectives.SymmetricArena`` (buffer / signal-pad pointers of every rank, multicast
    pointer, status words), built on the library's own driver-API bring-up."""

    flavour = "driver"

    def __init__(self, nbytes: int, device: torch.device, group=None, multicast: bool = True) -> None:
        self.group = group if group is not None else dist.group.WORLD
        self.device = torch.device(device)
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        if self.world > 8:
            raise RuntimeError("the peer-memory collectives cover one NVSwitch domain (<= 8 GPUs)")
        L = lib()
        self._declare(L)
        nbytes = (int(nbytes) + 4095) // 4096 * 4096
        self.nbytes = nbytes
        sig_bytes = SIGNAL_WORDS * 4
        self._h = C.c_void_p()
        size = C.c_longlong()
        with torch.cuda.device(self.device):
            check(L.b200dt_arena_create(self.device.index, self.rank, self.world, nbytes + sig_bytes, C.byref(self._h),
                                        C.byref(size)), "arena_create")
        self.size = int(size.value)
        # the run tag makes the abstract socket names unique per arena instance
        tag_box = [uuid.uuid4().hex[:12] if self.rank == 0 else None]
        dist.broadcast_object_list(tag_box, src=dist.get_global_rank(self.group, 0), group=self.group)
        tag = tag_box[0]
        fd = C.c_int(-1)
        check(L.b200dt_arena_export_fd(self._h, C.byref(fd)), "arena_export_fd")
        fds = _exchange_fds(self.group, self.rank, self.world, fd.value, tag + "m", device=self.device)
        try:
            for r in range(self.world):
                if r != self.rank:
                    rc = L.b200dt_arena_map_peer(self._h, r, fds[r])
                    if rc:
                        raise RuntimeError(f"arena_map_peer({r}) failed: rc {rc}, CUresult {L.b200dt_arena_last_error(self._h)}")
        finally:
            for f in fds:
                if f >= 0:
                    os.close(f)
        self.multicast_ptr = 0
        self.multicast_error = None
        if multicast and self.world > 1:
            self._bind_multicast(L, tag)
        base = [int(L.b200dt_arena_ptr(self._h, r)) for r in range(self.world)]
        self.buf_ptrs = base
        self.sig_ptrs = [b + nbytes for b in base]
        self._bufs_c = u64_array(self.buf_ptrs)
        self._sigs_c = u64_array(self.sig_ptrs)
        whole = torch.as_tensor(_Foreign(base[self.rank], nbytes + sig_bytes, self), device=self.device)
        if whole.data_ptr() != base[self.rank]:
            raise RuntimeError("torch copied the arena instead of aliasing it (cuda array interface)")
        self.buf = whole[:nbytes]
        self.sig = whole[nbytes:nbytes + sig_bytes].view(torch.int32)
        from .collectives import StatusWords
        self.status_words = StatusWords()
        torch.cuda.synchronize(self.device)
        _barrier(self.group, self.device)

    @staticmethod
    def _declare(L) -> None:
        h = L._handle
        if getattr(h, "_b200_arena_declared", False):
            return
        vp, i32, i64 = C.c_void_p, C.c_int, C.c_longlong
        h.b200dt_multicast_supported.argtypes = [i32]
        h.b200dt_arena_create.argtypes = [i32, i32, i32, i64, C.POINTER(vp), C.POINTER(i64)]
        h.b200dt_arena_export_fd.argtypes = [vp, C.POINTER(i32)]
        h.b200dt_arena_map_peer.argtypes = [vp, i32, i32]
        h.b200dt_arena_mc_create.argtypes = [vp, C.POINTER(i32)]
        h.b200dt_arena_mc_import.argtypes = [vp, i32]
        h.b200dt_arena_mc_add_device.argtypes = [vp]
        h.b200dt_arena_mc_bind.argtypes = [vp]
        h.b200dt_arena_ptr.argtypes = [vp, i32]
        h.b200dt_arena_ptr.restype = C.c_ulonglong
        h.b200dt_arena_mc_ptr.argtypes = [vp]
        h.b200dt_arena_mc_ptr.restype = C.c_ulonglong
        h.b200dt_arena_last_error.argtypes = [vp]
        h.b200dt_arena_destroy.argtypes = [vp]
        h._b200_arena_declared = True

    def _bind_multicast(self, L, tag: str) -> None:
        """All-or-nothing across the group: every step is followed by an agreement round, so either every
        rank ends up with a multicast address or none does (and the reason is kept in ``multicast_error``)."""
        def agree(ok: bool, why: str) -> bool:
            votes = [None] * self.world
            dist.all_gather_object(votes, (bool(ok), why), group=self.group)
            bad = [f"rank {r}: {w}" for r, (o, w) in enumerate(votes) if not o]
            if bad:
                self.multicast_error = "; ".join(bad)
            return not bad

        sup = int(L.b200dt_multicast_supported(self.device.index))
        if not agree(sup == 1, f"CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED={sup}"):
            return
        fd = C.c_int(-1)
        rc = 0
        if self.rank == 0:
            rc = L.b200dt_arena_mc_create(self._h, C.byref(fd))
        if not agree(rc == 0, f"cuMulticastCreate rc {rc} CUresult {L.b200dt_arena_last_error(self._h)}"):
            return
        fds = _exchange_fds(self.group, self.rank, self.world, fd.value, tag + "c", only_from=0, device=self.device)
        rc = 0
        if self.rank != 0:
            rc = L.b200dt_arena_mc_import(self._h, fds[0])
        for r, f in enumerate(fds):
            if f >= 0 and (r == 0):
                os.close(f)
        if not agree(rc == 0, f"multicast import rc {rc} CUresult {L.b200dt_arena_last_error(self._h)}"):
            return
        rc = L.b200dt_arena_mc_add_device(self._h)
        if not agree(rc == 0, f"cuMulticastAddDevice rc {rc} CUresult {L.b200dt_arena_last_error(self._h)}"):
            return                                   # (the agreement round is the barrier AddDevice needs)
        rc = L.b200dt_arena_mc_bind(self._h)
        if not agree(rc == 0, f"cuMulticastBindMem rc {rc} CUresult {L.b200dt_arena_last_error(self._h)}"):
            return
        self.multicast_ptr = int(L.b200dt_arena_mc_ptr(self._h))

    @property
    def has_multicast(self) -> bool:
        return self.multicast_ptr != 0

    @property
    def status(self):
        return self.status_words

    def check_status(self) -> None:
        self.status_words.check(self.rank)

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self.buf = self.sig = None
            lib().b200dt_arena_destroy(self._h)
            self._h = None
