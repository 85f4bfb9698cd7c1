"""Process-group bootstrap (control plane).

``init()`` / ``destroy()`` play the role of ``init_process_group`` / ``destroy_process_group`` in
the reference (/root/reference/ddp_gpus.py:12-17,93 and ddp_gpus_torchrun.py:12-14,88).  The
rendezvous (TCPStore), barriers and tiny metadata exchanges stay on ``torch.distributed``
(NCCL on GPUs, gloo on CPU); the gradient traffic goes through this library's own kernels over a
symmetric-memory arena that is attached to the group lazily (see ``ops/collectives.py``).
"""
from __future__ import annotations

import datetime
import os
import socket
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class GroupState:
    backend: str
    rank: int
    world_size: int
    local_rank: int
    device: torch.device
    arena: object = None            # ops.collectives.SymmetricArena once created
    extras: dict = field(default_factory=dict)


_state: Optional[GroupState] = None


def find_free_port() -> int:
    """A TCP port that is free right now (the reference hard-codes 12345, ddp_gpus.py:14)."""
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def default_backend() -> str:
    return "nccl" if torch.cuda.is_available() else "gloo"


def init(backend: Optional[str] = None, rank: Optional[int] = None, world_size: Optional[int] = None,
         master_addr: Optional[str] = None, master_port: Optional[int] = None,
         local_rank: Optional[int] = None, timeout_s: float = 600.0) -> GroupState:
    """Join the job.

    With no arguments everything comes from the torchrun environment contract (``RANK``,
    ``WORLD_SIZE``, ``LOCAL_RANK``, ``MASTER_ADDR``, ``MASTER_PORT``); explicit ``rank`` /
    ``world_size`` give the ``mp.spawn`` flavour where the parent hands each child its rank.
    On CUDA machines the process is pinned to ``cuda:local_rank`` *before* any communicator is
    created, so every later allocation and kernel lands on the right GPU.
    """
    global _state
    if _state is not None:
        raise RuntimeError("process group already initialised; call destroy() first")
    backend = backend or default_backend()
    if master_addr is not None:
        os.environ["MASTER_ADDR"] = str(master_addr)
    if master_port is not None:
        os.environ["MASTER_PORT"] = str(master_port)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if rank is None:
        rank = int(os.environ.get("RANK", "0"))
    if world_size is None:
        world_size = int(os.environ.get("WORLD_SIZE", "1"))
    auto_port = False
    if "MASTER_PORT" not in os.environ:
        if world_size == 1:
            os.environ["MASTER_PORT"] = str(find_free_port())
            auto_port = True
        else:
            raise RuntimeError("MASTER_PORT is not set (use the launcher or pass master_port=)")
    if local_rank is None:
        local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if backend == "nccl":
        if not torch.cuda.is_available():
            raise RuntimeError("backend 'nccl' needs CUDA; use backend='gloo' on CPU")
        n = torch.cuda.device_count()
        device = torch.device("cuda", local_rank % max(n, 1))
        torch.cuda.set_device(device)
    else:
        device = torch.device("cpu")
    kwargs = {}
    if backend == "nccl":
        kwargs["device_id"] = device
    for attempt in range(4):
        try:
            dist.init_process_group(backend=backend, rank=rank, world_size=world_size,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
            break
        except Exception as e:          # noqa: BLE001 -- c10d raises DistNetworkError / RuntimeError by version
            # a port this process picked itself can be taken between the probe and the bind: pick another one
            if not auto_port or attempt == 3 or "address already in use" not in str(e).lower():
                raise
            os.environ["MASTER_PORT"] = str(find_free_port())
    _state = GroupState(backend=backend, rank=rank, world_size=world_size, local_rank=local_rank,
                        device=device)
    return _state


def destroy() -> None:
    global _state
    if _state is None:
        return
    arena = _state.arena
    _state = None
    if arena is not None:
        try:
            arena.close()
        except Exception:
            pass
    if dist.is_initialized():
        dist.destroy_process_group()


def is_initialized() -> bool:
    return _state is not None


def state() -> GroupState:
    if _state is None:
        if dist.is_initialized():   # initialised by plain torch: adopt it
            return adopt()
        raise RuntimeError("process group is not initialised; call init() first")
    return _state


def adopt() -> GroupState:
    """Adopt a process group that the user created with plain ``torch.distributed``."""
    global _state
    if _state is not None:
        return _state
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised")
    backend = dist.get_backend()
    r, w = dist.get_rank(), dist.get_world_size()
    lr = int(os.environ.get("LOCAL_RANK", str(r)))
    if backend == "nccl":
        device = torch.device("cuda", torch.cuda.current_device())
    else:
        device = torch.device("cpu")
    _state = GroupState(backend=backend, rank=r, world_size=w, local_rank=lr, device=device)
    return _state


def rank(require_init: bool = False) -> int:
    if _state is None and not dist.is_initialized():
        if require_init:
            raise RuntimeError("Requires distributed package to be initialised (call init())")
        return 0
    return state().rank


def world_size(require_init: bool = False) -> int:
    if _state is None and not dist.is_initialized():
        if require_init:
            raise RuntimeError("Requires distributed package to be initialised (call init())")
        return 1
    return state().world_size


def barrier() -> None:
    st = state()
    if st.world_size > 1:
        if st.backend == "nccl":
            dist.barrier(device_ids=[st.device.index])
        else:
            dist.barrier()
