"""ctypes binding of ``_native/libb200dt.so`` (the C ABI declared in ``include/b200dt.h``).

The library is loaded lazily on first use.  On a machine with a GPU a missing or unloadable
library is a hard error: there is no eager / PyTorch fallback for the CUDA ops of this package.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

import torch

_lock = threading.Lock()
_lib: Optional["_Counted"] = None

# Number of calls into the kernel library (each is one or more kernel launches of THIS library);
# bench.py reports the per-step delta as "gpu_launches".
launch_count = 0


class _Counted:
    """Attribute proxy over the CDLL handle that counts kernel-launching calls."""

    def __init__(self, handle: C.CDLL, counted) -> None:
        self._handle = handle
        for name in counted:
            setattr(self, name, self._wrap(getattr(handle, name)))

    @staticmethod
    def _wrap(fn):
        def call(*args):
            global launch_count
            launch_count += 1
            return fn(*args)
        return call

    def __getattr__(self, name):
        return getattr(self._handle, name)

vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_uint64

_SIGS = {
    "b200dt_gemm_tn_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, i32, i32, vp, vp],
    "b200dt_gemm_nn_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp],
    "b200dt_gemm_nt_reduce_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp],
    "b200dt_conv2d_fprop_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp],
    "b200dt_conv2d_wgrad_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp],
    "b200dt_conv2d_dgrad_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp],
    "b200dt_conv2d_dgrad_acc_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp],
    "b200dt_stem_im2col": [vp, vp, i32, i32, i32, i32, vp],
    "b200dt_stem_conv_fprop": [vp, vp, vp, i32, i32, i32, vp, vp],
    "b200dt_stem_conv_wgrad": [vp, vp, vp, vp, i32, i32, i32, vp],
    "b200dt_cast_f32_to_bf16": [vp, vp, i64, vp],
    "b200dt_transpose_bf16": [vp, vp, i32, i32, vp],
    "b200dt_nchw_f32_to_nhwc_bf16": [vp, vp, i32, i32, i32, i32, vp],
    "b200dt_bn_stats": [vp, vp, i64, i32, vp],
    "b200dt_bn_apply": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp, vp, vp],
    "b200dt_relu_bwd_bias": [vp, vp, vp, vp, i64, i32, vp],
    "b200dt_skinny_linear_fwd": [vp, vp, vp, vp, i32, i32, i32, vp],
    "b200dt_skinny_linear_bwd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "b200dt_skinny_linear_fwd_host": [vp, vp, vp, vp, i32, i32, i32],
    "b200dt_skinny_linear_bwd_host": [vp, vp, vp, vp, vp, vp, i32, i32, i32],
    "b200dt_bn_apply_eval": [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, vp],
    "b200dt_bn_bwd_reduce": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp],
    "b200dt_bn_bwd_apply": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, vp],
    "b200dt_maxpool3x3s2_fwd": [vp, vp, vp, i32, i32, i32, i32, vp],
    "b200dt_maxpool3x3s2_bwd": [vp, vp, vp, i32, i32, i32, i32, vp],
    "b200dt_avgpool_fwd": [vp, vp, i32, i32, i32, vp],
    "b200dt_avgpool_bwd": [vp, vp, i32, i32, i32, vp],
    "b200dt_mse_loss": [vp, vp, vp, vp, i64, f32, i32, vp, vp],
    "b200dt_sum_loss": [vp, vp, vp, i64, f32, i32, vp, vp],
    "b200dt_ce_prob_loss": [vp, vp, vp, vp, i32, i32, f32, i32, vp, vp],
    "b200dt_ce_index_loss": [vp, vp, vp, vp, i32, i32, f32, i32, vp, vp],
    "b200dt_multi_sgd": [vp, vp, i32, f32, f32, f32, i32, vp],
    "b200dt_multi_adam": [vp, vp, i32, f32, f32, f32, f32, f32, i32, f32, i32, vp],
    "b200dt_multi_adam_capturable": [vp, vp, i32, f32, f32, f32, f32, f32, vp, f32, i32, vp],
    "b200dt_coll_launch": [i32, vp, vp, u64, i32, i32, i32, vp, vp, i32, i64, i32, i32, f32, i32,
                           i32, vp, i32, vp],
    "b200dt_scale_cast": [vp, vp, i64, f32, i32, vp],
    "b200dt_cast_to_f32": [vp, vp, i64, i32, vp],
    "b200dt_allreduce_local_emulate": [vp, i32, vp, i64, i32, f32, vp],
    "b200dt_enable_peer_access": [i32, i32],
    "b200dt_p2p_broadcast": [vp, vp, i32, i64, vp],
    "b200dt_p2p_reduce_f32": [vp, vp, i32, i64, f32, i32, vp],
    "b200dt_p2p_gather_rows": [vp, vp, i64, vp, i32, vp],
    "b200dt_p2p_multi_copy": [vp, vp, vp, i32, vp],
    "b200dt_memcpy_h2d_async": [vp, vp, i64, vp],
}


class NativeLibraryError(RuntimeError):
    pass


def available() -> bool:
    """True if the compiled kernel library exists on disk (it is only usable with a GPU)."""
    from .build import lib_path

    return lib_path().exists()


def lib() -> "_Counted":
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        from .build import lib_path

        path = lib_path()
        if not path.exists():
            raise NativeLibraryError(
                f"{path} is missing: run `python -m pytorch_distributed_training_tutorials_b200.build` "
                "(nvcc, sm_100a). The CUDA ops of this package have no PyTorch fallback."
            )
        try:
            handle = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        except OSError as e:  # pragma: no cover - depends on the box
            raise NativeLibraryError(f"cannot load {path}: {e}") from e
        for name, argtypes in _SIGS.items():
            fn = getattr(handle, name)
            fn.argtypes = argtypes
            fn.restype = i32
        handle.b200dt_loss_scratch_bytes.restype = i64
        handle.b200dt_loss_scratch_bytes.argtypes = []
        handle.b200dt_stem_wgrad_scratch_floats.restype = i64
        handle.b200dt_stem_wgrad_scratch_floats.argtypes = []
        handle.b200dt_coll_chunk_bytes.restype = i32
        handle.b200dt_coll_chunk_bytes.argtypes = []
        handle.b200dt_device_sm_count.restype = i32
        handle.b200dt_set_deterministic.restype = None
        handle.b200dt_set_deterministic.argtypes = [i32]
        handle.b200dt_get_deterministic.restype = i32
        handle.b200dt_get_deterministic.argtypes = []
        handle.b200dt_bn_replicas.restype = i32
        handle.b200dt_bn_replicas.argtypes = []
        handle.b200dt_bn_replicas_used.restype = i32
        handle.b200dt_bn_replicas_used.argtypes = [i32]
        handle.b200dt_set_sums_prezeroed.restype = None
        handle.b200dt_set_sums_prezeroed.argtypes = [i32]
        handle.b200dt_get_sums_prezeroed.restype = i32
        handle.b200dt_get_sums_prezeroed.argtypes = []
        handle.b200dt_capture_id.restype = C.c_ulonglong
        handle.b200dt_capture_id.argtypes = [C.c_void_p]
        handle.b200dt_version.restype = C.c_char_p
        handle.b200dt_error_string.restype = C.c_char_p
        handle.b200dt_error_string.argtypes = [i32]
        _lib = _Counted(handle, [n for n in _SIGS if n not in ("b200dt_enable_peer_access", "b200dt_memcpy_h2d_async")
                                 and not n.endswith("_host")])
        return _lib


_mode_listeners: list = []      # callbacks run when a library-wide mode changes (caches keyed on it)


def set_deterministic(on: bool = True) -> None:
    """Fixed-order reductions in every kernel of the library (BatchNorm statistics, split-K weight gradients,
    bias gradients): the same step on the same inputs gives the same bits.  Also settable with
    ``B200DT_DETERMINISTIC=1`` in the environment.  Slower; meant for regression tests and debugging."""
    lib().b200dt_set_deterministic(int(bool(on)))
    for fn in _mode_listeners:
        fn()


def is_deterministic() -> bool:
    return bool(lib().b200dt_get_deterministic())


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    if rc > 0:
        try:
            msg = lib().b200dt_error_string(rc).decode()
        except Exception:  # pragma: no cover
            msg = "cuda error"
        raise RuntimeError(f"b200dt {what}: CUDA error {rc}: {msg}")
    reasons = {
        -1: "cuTensorMapEncodeTiled entry point not found",
        -2: "tile configuration does not fit in shared memory",
        -3: "unsupported shape / alignment / dtype for this kernel",
        -4: "NVLS requested but no multicast address",
        -5: "not enough signal-pad words for this grid",
        -6: "peer access not possible between these devices",
        -7: "a CUDA driver call failed (b200dt_arena_last_error has the CUresult)",
        -9: "input larger than the fixed scratch of this entry point",
    }
    raise RuntimeError(f"b200dt {what}: error {rc}: {reasons.get(rc, 'tensor map encode failed' if rc < -1000 else 'invalid argument')}")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device=None) -> int:
    """Raw cudaStream_t of the current stream of ``device`` (fast path: no Stream object)."""
    if _raw_stream is not None:
        if device is None:
            idx = torch.cuda.current_device()
        elif isinstance(device, int):
            idx = device
        else:
            idx = device.index if device.index is not None else torch.cuda.current_device()
        return _raw_stream(idx)
    return torch.cuda.current_stream(device).cuda_stream


class on_device:
    """``with on_device(dev):`` like ``torch.cuda.device`` but free when ``dev`` is already current
    (the common case: one process per GPU), which matters at ~500 kernel launches per step."""

    __slots__ = ("idx", "prev")

    def __init__(self, device) -> None:
        self.idx = device.index if isinstance(device, torch.device) else device
        self.prev = -1

    def __enter__(self):
        if self.idx is not None:
            cur = torch.cuda.current_device()
            if cur != self.idx:
                self.prev = cur
                torch.cuda.set_device(self.idx)
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch.cuda.set_device(self.prev)
            self.prev = -1
        return False


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def u64_array(values) -> C.Array:
    arr = (u64 * len(values))(*[int(v) for v in values])
    return arr


class PinnedUploader:
    """Stream-ordered upload of small host tables through a ring of pinned staging buffers.

    Outside graph capture a slot is reused only after the event recorded behind its last copy has
    completed.  During capture no event is recorded (it could never be queried afterwards); the
    slot is frozen instead, because the captured memcpy node re-reads it on every replay."""

    def __init__(self, slots: int = 4) -> None:
        self._slots = []          # [buffer(uint8, pinned), event or None, frozen]
        self._max_free = slots

    def _acquire(self, nbytes: int, capturing: bool):
        if capturing:
            # no event may be queried while a capture is in progress: take a fresh, dedicated slot
            s = [torch.empty(max(nbytes, 4096), dtype=torch.uint8).pin_memory(), None, True]
            self._slots.append(s)
            return s
        for s in self._slots:
            if s[2] or s[0].numel() < nbytes:
                continue
            if s[1] is None or s[1].query():
                return s
        free = [s for s in self._slots if not s[2] and s[0].numel() >= nbytes]
        if len(free) >= self._max_free:
            free[0][1].synchronize()
            return free[0]
        s = [torch.empty(max(nbytes, 4096), dtype=torch.uint8).pin_memory(), None, False]
        self._slots.append(s)
        return s

    def upload(self, dst: torch.Tensor, src_cpu: torch.Tensor, stream: "torch.cuda.Stream") -> None:
        nbytes = src_cpu.numel() * src_cpu.element_size()
        assert dst.numel() * dst.element_size() >= nbytes
        capturing = torch.cuda.is_current_stream_capturing()
        slot = self._acquire(nbytes, capturing)
        slot[0][:nbytes].copy_(src_cpu.contiguous().view(torch.uint8).view(-1))
        check(lib().b200dt_memcpy_h2d_async(dst.data_ptr(), slot[0].data_ptr(), nbytes, stream.cuda_stream),
              "table upload")
        if capturing:
            slot[2] = True
        else:
            if slot[1] is None:
                slot[1] = torch.cuda.Event()
            slot[1].record(stream)
