"""``DistributedDataParallel``: multi-process data parallelism, one process per GPU.

Same construction-time and per-step contract as the wrapper the reference uses at
/root/reference/ddp_gpus.py:32 (``DDP(model, device_ids=[gpu_id])``, all defaults):

* ctor: parameter shapes are verified across ranks, then rank 0's parameters and buffers are
  broadcast so that every replica starts identical;
* forward: buffers (BatchNorm running statistics) are re-broadcast from rank 0 when
  ``broadcast_buffers`` is set, the wrapped module runs, and the C++ reducer is armed;
* backward: one autograd post-hook per parameter; when the last gradient of a bucket is ready the
  bucket is reduced *while backward keeps running*: on GPUs by ONE fused sm_100a kernel on a side
  stream (gather grads -> 1/world scale -> wire-dtype cast -> one-shot / two-shot / NVLS reduce
  over NVLink peer memory -> result written straight into ``.grad``), on CPU / gloo by
  ``dist.all_reduce`` on a flat bucket;
* first iteration uses one bucket holding everything, then buckets are rebuilt in the observed
  gradient-ready order with the 1 MiB / ``bucket_cap_mb`` caps, rank 0's layout being authoritative;
* ``no_sync()`` skips the reduction for gradient accumulation; ``state_dict()`` keys carry the
  ``module.`` prefix exactly like torch's wrapper.
"""
from __future__ import annotations

import contextlib
import ctypes
import os
import sys
import time
from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from .. import group as _group
from ..utils.timing import TRACE as _TRACE, trace_range
from ._native_reducer import load as _load_reducer

FIRST_BUCKET_BYTES = 1024 * 1024


class _Bucket:
    def __init__(self, index: int, param_ids: List[int]) -> None:
        self.index = index
        self.param_ids = param_ids
        self.table = None          # ops.collectives.TensorTable (GPU path)
        self.flat = None           # flat buffer (fallback path)
        self.work = None
        self.nbytes = 0
        self.mode = None
        self.stage_base = 0          # byte offset of this bucket's region inside the arena
        self.views = None            # gradient views living in that region (gradient_as_bucket_view)


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids: Optional[List[int]] = None, output_device=None,
                 broadcast_buffers: bool = True, process_group=None, bucket_cap_mb: float = 25.0,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = False,
                 init_sync: bool = True, allreduce: str = "auto", comm_dtype: str = "fp32",
                 max_comm_blocks: int = 32) -> None:
        super().__init__()
        if find_unused_parameters:
            raise NotImplementedError("find_unused_parameters=True is not supported (the reference never uses it)")
        if not dist.is_initialized():
            raise RuntimeError("Default process group has not been initialized, please make sure to call "
                               "init() / init_process_group.")
        self.module = module
        self.process_group = process_group if process_group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.process_group)
        self.rank = dist.get_rank(self.process_group)
        self.broadcast_buffers = broadcast_buffers
        self.bucket_bytes_cap = int(bucket_cap_mb * 1024 * 1024)
        self.require_backward_grad_sync = True
        self.allreduce = allreduce
        self.comm_dtype = comm_dtype
        self.max_comm_blocks = max_comm_blocks
        self.gradient_as_bucket_view = gradient_as_bucket_view
        # B200DT_DEBUG=1: before every bucket collective all ranks compare (sequence number, bucket
        # index, byte size); a rank that issues collectives in another order raises instead of
        # hanging in a flag barrier (the role ProcessGroupWrapper plays for torch's DDP)
        self._debug = os.environ.get("B200DT_DEBUG", "0") not in ("", "0")
        self._debug_seq = 0

        self._params = [p for p in module.parameters() if p.requires_grad]
        if not self._params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter "
                               "that requires a gradient.")
        devs = {p.device for p in self._params}
        if len(devs) != 1:
            raise ValueError("DistributedDataParallel expects all parameters on one device "
                             f"(found {sorted(map(str, devs))}); multi-device modules belong to parallel.model_parallel")
        self.device = next(iter(devs))
        if device_ids is not None:
            if len(device_ids) != 1:
                raise ValueError("device_ids can only be None or contain a single element.")
            d = device_ids[0]
            d = torch.device("cuda", d) if isinstance(d, int) else torch.device(d)
            if d != self.device:
                raise ValueError(f"module parameters live on {self.device} but device_ids={device_ids}")
        self.device_ids = device_ids
        self.output_device = output_device
        self._use_kernels = (self.device.type == "cuda" and self.world_size > 1 and allreduce != "nccl"
                             and all(p.dtype in (torch.float32, torch.bfloat16) for p in self._params))
        self._arena = None
        self._side_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._ready_event = torch.cuda.Event() if self.device.type == "cuda" else None
        self._fallback_reason = None

        if init_sync and self.world_size > 1:
            self._verify_params_across_ranks()
        if self._use_kernels:
            self._create_arena()
        if init_sync and self.world_size > 1:
            self._sync_module_states()

        R = _load_reducer()
        self._R = R
        # iteration 1: one bucket holding every parameter of a dtype (launch once, learn the ready order)
        self._rebuilt = False
        first, _ = R.assign_buckets(self._params, [sys.maxsize])
        first = [list(ids) for ids in first]
        self._buckets = self._make_buckets(first)
        self._reducer = R.Reducer(self._params, first, self._on_bucket_ready, self._on_finalize)
        self._reducer.set_record_order(True)
        self._stats = {"allreduce_bytes": 0, "launches": 0, "rebuilds": 0}
        self._install_native_plan()
        self._t_ctor = time.time()

    # ------------------------------------------------------------------ construction helpers
    def _verify_params_across_ranks(self) -> None:
        """Every rank must hold the same number of parameters with the same shapes (X1)."""
        desc = [tuple(p.shape) for p in self._params]
        h = hash(tuple(desc)) & 0x7FFFFFFFFFFFFFFF
        mine = torch.tensor([len(desc), h, sum(p.numel() for p in self._params)], dtype=torch.int64,
                            device=self.device if self.device.type == "cuda" else "cpu")
        allv = [torch.zeros_like(mine) for _ in range(self.world_size)]
        dist.all_gather(allv, mine, group=self.process_group)
        for r, v in enumerate(allv):
            if not torch.equal(v.cpu(), mine.cpu()):
                raise RuntimeError(
                    f"DDP expects same model across all ranks, but Rank {self.rank} has {int(mine[0])} params "
                    f"({int(mine[2])} elements), while rank {r} has {int(v[0])} params ({int(v[2])} elements).")

    def _create_arena(self) -> None:
        from ..ops.collectives import make_arena
        st = _group.state() if _group.is_initialized() or dist.is_initialized() else None
        total = sum((p.numel() + 7) // 8 * 8 * 4 for p in self._params)
        # the construction-time broadcast stages parameters AND buffers (as 32-bit words) at once
        bufs = sum((b.numel() * b.element_size() // 4 + 8) // 8 * 8 * 4 for b in self.module.buffers())
        need = total + bufs + (1 << 20)
        # gradients that LIVE in the arena get a region of their own behind the broadcast staging
        self._grad_region = (need + 4095) // 4096 * 4096
        if self.gradient_as_bucket_view:
            need = self._grad_region + total + 256 * (len(self._params) + 64)
        try:
            if not self.gradient_as_bucket_view and st is not None and st.arena is not None \
                    and st.arena.nbytes >= need and st.arena.group is self.process_group:
                self._arena = st.arena
            else:
                with torch.cuda.device(self.device):
                    self._arena = make_arena(need, self.device, self.process_group)
                if st is not None and st.arena is None and not self.gradient_as_bucket_view:
                    st.arena = self._arena
        except Exception as e:  # symmetric memory unavailable: say so loudly, fall back to NCCL
            self._use_kernels = False
            self._fallback_reason = f"{type(e).__name__}: {e}"
            print(f"[b200dt] WARNING rank {self.rank}: symmetric-memory arena unavailable ({self._fallback_reason}); "
                  "gradient all-reduce falls back to NCCL", file=sys.stderr, flush=True)
        flag = torch.tensor([1 if self._use_kernels else 0], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.process_group)
        if int(flag.item()) == 0:
            self._use_kernels = False

    def _broadcast_tensors(self, tensors: List[torch.Tensor]) -> None:
        """Rank 0's values overwrite everybody's. On GPUs every tensor whose byte size is a multiple
        of 4 rides ONE peer-store kernel as raw 32-bit words (bit-exact for any dtype)."""
        if not tensors:
            return
        fast, slow = [], []
        for t in tensors:
            if (self._use_kernels and t.is_cuda and t.is_contiguous() and t.numel() > 0
                    and (t.numel() * t.element_size()) % 4 == 0 and t.data_ptr() % 4 == 0):
                fast.append(t.view(-1).view(torch.int32) if t.dtype != torch.float32 else t.view(-1))
            elif t.numel() > 0:
                slow.append(t)
        if fast:
            from ..ops import collectives as coll
            cur = torch.cuda.current_stream(self.device)
            key = tuple(t.numel() for t in fast)
            tables = self.__dict__.setdefault("_bcast_tables", {})
            table = tables.get(key)
            if table is None:
                table = coll.TensorTable(list(key), self.device, False, False)
                tables[key] = table
            table.bind(fast, cur)
            coll.launch("bcast", self._arena, table, 1.0, cur, root=0,
                        sig_base=coll.signal_base(self._arena, f"ddp{id(self)}:bcast", self.max_comm_blocks),
                        max_blocks=self.max_comm_blocks)
        if slow:
            src = dist.get_global_rank(self.process_group, 0)
            by_dtype = {}
            for t in slow:
                by_dtype.setdefault((t.dtype, t.device), []).append(t)
            for ts in by_dtype.values():          # coalesced: one broadcast per dtype
                flat = torch.cat([t.reshape(-1) for t in ts])
                dist.broadcast(flat, src=src, group=self.process_group)
                off = 0
                for t in ts:
                    n = t.numel()
                    t.copy_(flat[off:off + n].view_as(t))
                    off += n

    def _sync_module_states(self) -> None:
        """Rank 0's parameters and buffers overwrite everybody else's (X2)."""
        with torch.no_grad():
            self._broadcast_tensors([p.data for p in self.module.parameters()] + [b.data for b in self.module.buffers()])

    def _sync_buffers(self) -> None:
        """Per-forward buffer broadcast from rank 0 (X3)."""
        bufs = [b.data for b in self.module.buffers()]
        if bufs:
            with torch.no_grad():
                self._broadcast_tensors(bufs)

    def _make_buckets(self, layout: List[List[int]]) -> List[_Bucket]:
        out = []
        for i, ids in enumerate(layout):
            b = _Bucket(i, list(ids))
            b.nbytes = sum(self._params[j].numel() * self._params[j].element_size() for j in ids)
            if self._use_kernels:
                from ..ops import collectives as coll
                tb16 = self._params[ids[0]].dtype == torch.bfloat16
                b.table = coll.TensorTable([self._params[j].numel() for j in ids], self.device,
                                           self.comm_dtype == "bf16", tb16, target_blocks=self.max_comm_blocks)
                b.mode = coll.pick_mode(b.table.stage_bytes, self._arena, self.allreduce, b.table.wire_bf16)
            out.append(b)
        self._assign_grad_views(out)
        return out

    def _install_native_plan(self) -> None:
        """Hand the C++ reducer everything it needs to launch the fused collective of each bucket itself
        (no GIL, no Python in the backward hot path).  The Python callbacks stay in charge on CPU / gloo, with
        the NCCL opt-out, and when the debug handshake or NVTX tracing wrap every launch."""
        if (not self._use_kernels or self.world_size == 1 or self._debug or _TRACE
                or os.environ.get("B200DT_PY_REDUCER", "0") not in ("", "0")):
            self._reducer.clear_native_plan()
            return
        from ..ops import collectives as coll
        from .._lib import lib
        a = self._arena
        sig = coll.signal_base(a, f"ddp{id(self)}:buckets", self.max_comm_blocks)
        self._bucket_sig_base = sig
        buckets = []
        for b in self._buckets:
            t = b.table
            buckets.append(dict(mode=coll.MODE[b.mode], n_chunks=t.n_chunks, tensor_bf16=int(t.tensor_bf16),
                                wire_bf16=int(t.wire_bf16), sig_base=sig, grid=t.grid(self.max_comm_blocks),
                                in_place=int(b.views is not None), stage_base=int(b.stage_base),
                                scale=1.0 / self.world_size, tensors_dev=t.tensors.data_ptr(),
                                chunks_dev=t.chunks.data_ptr(), nbytes=int(b.nbytes), param_ids=list(b.param_ids),
                                numels=list(t.numels), views=list(b.views) if b.views is not None else None))
            if b.stage_base + t.stage_bytes > a.nbytes:
                raise RuntimeError(f"bucket {b.index} does not fit the arena")
        fn = ctypes.cast(lib()._handle.b200dt_coll_launch, ctypes.c_void_p).value
        self._reducer.set_native_plan(dict(
            launch_fn=fn, bufs=list(a.buf_ptrs), sigs=list(a.sig_ptrs), mc=int(a.multicast_ptr),
            status_dev=a.status_words.dev_ptr, status_host=ctypes.addressof(a.status_words._host.contents),
            rank=a.rank, world=a.world, device=self.device.index, side_stream=self._side_stream.cuda_stream,
            buckets=buckets))

    def _views_possible(self) -> bool:
        wire16 = self.comm_dtype == "bf16"
        return (self.gradient_as_bucket_view and self._use_kernels and self._arena is not None
                and all((p.dtype == torch.bfloat16) == wire16 for p in self._params))

    def _assign_grad_views(self, buckets) -> None:
        """gradient_as_bucket_view: every bucket owns a region of the symmetric arena and the
        parameters' ``.grad`` become views of it.  Weight-gradient kernels write there directly
        (``p._b200_grad_slot``), the all-reduce runs in place and nothing is copied back."""
        for p in self._params:
            if hasattr(p, "_b200_grad_slot"):
                p._b200_grad_slot = None
        if not self._views_possible():
            return
        from ..ops import collectives as coll
        base = self._grad_region
        for b in buckets:
            b.stage_base = base
            shapes = [self._params[j].shape for j in b.param_ids]
            b.views = coll.stage_views(self._arena, b.table, shapes, base)
            if b.mode == "oneshot":
                b.mode = "twoshot"            # in place, every rank may only overwrite the slice it owns
            region = self._arena.buf[base:base + b.table.stage_bytes].view(b.views[0].dtype)
            for k, j in enumerate(b.param_ids):
                p = self._params[j]
                p._b200_grad_slot = (region, b.table.tensor_offsets[k])
                if p.grad is not None:        # a stale view of the previous layout must not survive
                    p.grad = p.grad.detach().clone()
            base += (b.table.stage_bytes + 255) // 256 * 256

    def _rebuild_buckets(self) -> None:
        """Re-bucket in gradient-ready order with caps [1 MiB, bucket_cap]; rank 0 decides."""
        order = list(self._reducer.ready_order())
        if len(order) != len(self._params):
            return
        if self.world_size > 1:
            t = torch.tensor(order, dtype=torch.int64, device=self.device if self.device.type == "cuda" else "cpu")
            dist.broadcast(t, src=dist.get_global_rank(self.process_group, 0), group=self.process_group)
            order = t.cpu().tolist()
        tensors = [self._params[i] for i in order]
        layout, _ = self._R.assign_buckets(tensors, [FIRST_BUCKET_BYTES, self.bucket_bytes_cap], order)
        self._buckets = self._make_buckets(layout)
        self._reducer.set_buckets(layout)
        self._reducer.set_record_order(False)
        self._install_native_plan()
        self._rebuilt = True
        self._stats["rebuilds"] += 1

    def _debug_handshake(self, bi: int, b: "_Bucket") -> None:
        if self.device.type == "cuda" and torch.cuda.is_current_stream_capturing():
            return                                    # host round trips cannot be captured
        self._debug_seq += 1
        dev = self.device if dist.get_backend(self.process_group) == "nccl" else torch.device("cpu")
        mine = torch.tensor([self._debug_seq, bi, b.nbytes], dtype=torch.int64, device=dev)
        allv = [torch.zeros_like(mine) for _ in range(self.world_size)]
        dist.all_gather(allv, mine, group=self.process_group)
        for r, v in enumerate(allv):
            if not torch.equal(v.cpu(), mine.cpu()):
                raise RuntimeError(f"collective mismatch: rank {self.rank} is at (seq, bucket, bytes)={mine.tolist()} "
                                   f"but rank {r} is at {v.tolist()}")

    # ------------------------------------------------------------------ reducer callbacks
    def _on_bucket_ready(self, bi: int) -> None:
        b = self._buckets[bi]
        if self.world_size == 1:
            return
        grads = []
        for j in b.param_ids:
            g = self._params[j].grad
            if g is None:
                raise RuntimeError(f"parameter {j} has no gradient at bucket launch")
            if not g.is_contiguous():
                if self._use_kernels:
                    raise RuntimeError(f"parameter {j} has a non-contiguous .grad; the fused all-reduce reads gradients "
                                       "in place (make the gradient contiguous or use allreduce='nccl')")
                g = g.contiguous()          # fallback path: _on_finalize re-assigns p.grad from the flat bucket
            grads.append(g)
        self._stats["allreduce_bytes"] += b.nbytes
        self._stats["launches"] += 1
        if self._debug:
            self._debug_handshake(bi, b)
        if self._use_kernels:
            from ..ops import collectives as coll
            cur = torch.cuda.current_stream(self.device)
            self._ready_event.record(cur)
            self._side_stream.wait_event(self._ready_event)
            if not torch.cuda.is_current_stream_capturing():
                for g in grads:
                    g.record_stream(self._side_stream)
            b.table.bind(grads, self._side_stream)
            with trace_range(f"allreduce[{bi}]"):
                coll.launch(b.mode, self._arena, b.table, 1.0 / self.world_size, self._side_stream,
                            sig_base=coll.signal_base(self._arena, f"ddp{id(self)}:buckets", self.max_comm_blocks),
                            max_blocks=self.max_comm_blocks, stage_base=b.stage_base, in_place=b.views is not None)
        else:
            # gloo / NCCL path: flat bucket, pre-scale, async all-reduce (what torch's Reducer does)
            flat = torch.cat([g.reshape(-1) for g in grads])
            if self.comm_dtype == "bf16" and flat.dtype == torch.float32 and flat.is_cuda:
                flat = flat.to(torch.bfloat16)
            flat.div_(self.world_size)
            b.flat = flat
            b.grads = grads
            b.work = dist.all_reduce(flat, group=self.process_group, async_op=True)

    def _on_finalize(self) -> None:
        if self.world_size == 1:
            return
        if self._use_kernels:
            torch.cuda.current_stream(self.device).wait_stream(self._side_stream)
            self._arena.check_status()          # host-mapped word: no stream is touched
            for b in self._buckets:
                if b.views is None:
                    continue
                # the reduced gradients are in the arena: point .grad at them (no copy back)
                for view, j in zip(b.views, b.param_ids):
                    p = self._params[j]
                    if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                        p.grad = view
        else:
            for b in self._buckets:
                if b.work is None:
                    continue
                b.work.wait()
                off = 0
                for g, j in zip(b.grads, b.param_ids):
                    n = g.numel()
                    piece = b.flat[off:off + n].view_as(g)
                    p = self._params[j]
                    if p.grad is g:
                        g.copy_(piece)
                    else:
                        p.grad = piece.to(p.dtype).clone()
                    off += n
                b.work = None
                b.flat = None
                b.grads = None

    # ------------------------------------------------------------------ nn.Module API
    def forward(self, *inputs, **kwargs):
        with torch.autograd.profiler.record_function("b200dt.DistributedDataParallel.forward"):
            sync = torch.is_grad_enabled() and self.require_backward_grad_sync
            if sync and not self._rebuilt and self._reducer.finalized() and len(self._reducer.ready_order()) == len(self._params):
                self._rebuild_buckets()
            if self.world_size > 1 and self.broadcast_buffers and self.require_backward_grad_sync:
                self._sync_buffers()
            out = self.module(*inputs, **kwargs)
            if sync:
                self._reducer.prepare_for_backward()
            return out

    @contextlib.contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the context (gradients accumulate locally)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        self._reducer.set_require_sync(False)
        try:
            yield
        finally:
            self.require_backward_grad_sync = old
            self._reducer.set_require_sync(old)

    def rebind_autograd_hooks(self) -> None:
        """Re-create the parameters' gradient accumulators under the CURRENT stream (call inside a
        side-stream context before capturing the step in a CUDA graph)."""
        self._reducer.install_hooks()

    def get_logging_data(self) -> dict:
        d = dict(self._reducer.stats())
        d.update(self._stats)
        d["allreduce_bytes"] += d.get("native_bytes", 0)
        d["launches"] += d.get("native_launches", 0)
        d.update(world_size=self.world_size, rank=self.rank, rebuilt=self._rebuilt,
                 bucket_sizes_mib=[round(b.nbytes / 2**20, 2) for b in self._buckets],
                 bucket_modes=[b.mode for b in self._buckets],
                 comm="b200dt-peer-memory" if self._use_kernels else (dist.get_backend(self.process_group) if self.world_size > 1 else "none"),
                 comm_dtype=self.comm_dtype, fallback_reason=self._fallback_reason,
                 multicast=bool(self._arena.has_multicast) if self._arena is not None else False)
        return d

    def check_comm_status(self) -> None:
        """Raise ``CollectiveTimeout`` if a peer-memory collective gave up waiting for a peer.  Reads a
        host-mapped word (no synchronisation); the reducer's finalize does the same after every backward."""
        if self._arena is not None:
            self._arena.check_status()
