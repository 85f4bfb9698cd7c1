"""Tensor-level wrappers over the C ABI (one Python function per kernel entry point).

These do shape / dtype checks, allocate outputs with torch's caching allocator and launch on the
current CUDA stream.  No autograd here; see ``functional.py`` for the differentiable ops.
All activations are NHWC bf16 (``[N, H, W, C]`` contiguous), weights ``[Cout, R, S, Cin]`` bf16.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import _lib as _lib_mod
from .._lib import check, lib, on_device, ptr, stream_ptr

BF16 = torch.bfloat16
F32 = torch.float32


def bn_replicas() -> int:
    """Upper bound on the copies of the partial-sum array behind a BatchNorm ``sums`` buffer (16; 160 in
    deterministic mode); ``bn_replicas_used(C)`` is the exact count for a C-channel layer."""
    return int(lib().b200dt_bn_replicas())


_used_cache: dict = {}          # C -> copies; cleared when the deterministic switch changes (_lib.set_deterministic)
_lib_mod._mode_listeners.append(_used_cache.clear)


def bn_replicas_used(C: int) -> int:
    r = _used_cache.get(C)
    if r is None:
        r = _used_cache[C] = int(lib().b200dt_bn_replicas_used(int(C)))
    return r


# Zero-filled scratch for the statistics kernels.  Every conv-with-statistics, bn_stats and bn_bwd_reduce launch
# needs its partial-sum copies and ticket word zeroed first; as a memset per launch that is ~160 extra nodes in a
# ResNet-50 step (~3 us each when replayed from a graph).  Instead the buffers are carved, once each, from chunks
# that ONE fill zeroes, and the library is told not to memset (b200dt_set_sums_prezeroed).  A chunk belongs to one
# (device, stream, stream-capture) triple: the fill is ordered before every kernel that uses a slice, and a
# captured graph re-runs its own fill on every replay.
_ZERO_CHUNK_FLOATS = 1 << 20
_zero_chunks: dict = {}


_zero_fns = None          # (capture_id, get_prezeroed, set_prezeroed) bound once: this runs ~100x per step


def zeroed_floats(n: int, device) -> torch.Tensor:
    """``n`` fp32 zeros on ``device`` (256-byte aligned), valid for ONE accumulate-into use on the current stream."""
    if not isinstance(device, torch.device):
        device = torch.device(device)
    if device.type != "cuda":
        return torch.zeros(n, device=device, dtype=F32)
    global _zero_fns
    if _zero_fns is None:
        h = lib()._handle
        _zero_fns = (h.b200dt_capture_id, h.b200dt_get_sums_prezeroed, h.b200dt_set_sums_prezeroed)
        if not _zero_fns[1]():
            _zero_fns[2](1)
    idx = device.index
    if idx is None:
        idx = torch.cuda.current_device()
    st = stream_ptr(idx)
    cap = _zero_fns[0](st)
    n64 = (n + 63) & ~63
    ent = _zero_chunks.get((idx, st))
    if ent is None or ent[2] != cap or ent[1] + n64 > ent[3]:
        size = max(_ZERO_CHUNK_FLOATS, n64)
        with torch.cuda.device(idx):
            ent = [torch.zeros(size, device=device, dtype=F32), 0, cap, size]
        _zero_chunks[(idx, st)] = ent
    off = ent[1]
    ent[1] = off + n64
    return ent[0].narrow(0, off, n)


def _need(t: torch.Tensor, dtype, name: str) -> None:
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous CUDA {dtype} tensor, got {t.dtype} on {t.device} "
                         f"(contiguous={t.is_contiguous()})")


# --------------------------------------------------------------------------- GEMM / conv
def bn_sums_buffer(C: int, device) -> torch.Tensor:
    """Scratch for BatchNorm statistics: replicated {sum, sum of squares} + ticket (see bn.cu), zeroed.
    Good for ONE launch: the kernels add into it and do not clear it (``zeroed_floats``)."""
    return zeroed_floats(2 * C * bn_replicas_used(C) + 1, device)


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
            out_dtype=BF16, out: Optional[torch.Tensor] = None, bn_sums: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a[M,K] @ b[N,K]^T (+bias) (relu)``; bf16 inputs, fp32 accumulation on tcgen05.
    ``bn_sums``: optional buffer from ``bn_sums_buffer`` that receives the column statistics of the output."""
    _need(a, BF16, "a"); _need(b, BF16, "b")
    M, K = a.shape
    N, K2 = b.shape
    if K != K2 or K % 8:
        raise ValueError(f"gemm_tn: K mismatch or K % 8 != 0 ({K}, {K2})")
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    if bias is not None:
        _need(bias, F32, "bias")
    with on_device(a.device):
        check(lib().b200dt_gemm_tn_bf16(ptr(a), ptr(b), ptr(out), M, N, K, K, K, out.stride(0),
                                        ptr(bias), int(relu), int(out.dtype == F32), ptr(bn_sums),
                                        stream_ptr(a.device)), "gemm_tn")
    return out


def gemm_nn(a: torch.Tensor, b: torch.Tensor, out_dtype=BF16) -> torch.Tensor:
    """``a[M,K] @ b[K,N]`` with both operands row-major (b is read as an MN-major tcgen05 operand)."""
    _need(a, BF16, "a"); _need(b, BF16, "b")
    M, K = a.shape
    K2, N = b.shape
    if K != K2 or K % 8 or N % 8:
        raise ValueError("gemm_nn: K mismatch or K/N % 8 != 0")
    out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    with on_device(a.device):
        check(lib().b200dt_gemm_nn_bf16(ptr(a), ptr(b), ptr(out), M, N, K, K, N, out.stride(0),
                                        int(out_dtype == F32), stream_ptr(a.device)), "gemm_nn")
    return out


def gemm_nt_reduce(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a[R,M]^T @ b[R,N]`` -> fp32 ``[M,N]`` (weight gradients; reduction over rows)."""
    _need(a, BF16, "a"); _need(b, BF16, "b")
    R, M = a.shape
    R2, N = b.shape
    if R != R2 or M % 8 or N % 8:
        raise ValueError("gemm_nt_reduce: row mismatch or M/N % 8 != 0")
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=F32)
    with on_device(a.device):
        check(lib().b200dt_gemm_nt_reduce_bf16(ptr(a), ptr(b), ptr(out), M, N, R, M, N, out.stride(0), 0,
                                               stream_ptr(a.device)), "gemm_nt_reduce")
    return out


def relu_bwd_bias(dy: torch.Tensor, y: Optional[torch.Tensor], want_bias_grad: bool):
    """Backward glue of ``relu(linear)`` in one pass over ``[rows, C]`` bf16: returns
    ``(g, db)`` with ``g = dy * (y > 0)`` (``dy`` itself when ``y`` is None) and ``db`` the fp32 column sums."""
    _need(dy, BF16, "dy")
    rows, Cc = dy.shape
    g = dy
    if y is not None:
        _need(y, BF16, "y")
        g = torch.empty_like(dy)
    db = torch.empty(Cc, device=dy.device, dtype=F32) if want_bias_grad else None
    if y is None and db is None:
        return g, None
    with on_device(dy.device):
        check(lib().b200dt_relu_bwd_bias(ptr(dy), ptr(y), ptr(g) if y is not None else None, ptr(db), rows, Cc,
                                         stream_ptr(dy.device)), "relu_bwd_bias")
    return g, db


def skinny_linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """fp32 ``x[B,K] @ w[N,K]^T + bias`` on the SIMT kernel (any K, N; csrc/skinny.cu)."""
    _need(x, F32, "x"); _need(w, F32, "w")
    if bias is not None:
        _need(bias, F32, "bias")
    B, K = x.shape
    N = w.shape[0]
    y = torch.empty((B, N), device=x.device, dtype=F32)
    with on_device(x.device):
        check(lib().b200dt_skinny_linear_fwd(ptr(x), ptr(w), ptr(bias), ptr(y), B, K, N, stream_ptr(x.device)),
              "skinny_linear_fwd")
    return y


def skinny_linear_bwd(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, want_dx: bool, want_dw: bool,
                      want_db: bool):
    """Returns ``(dx, dw, db)`` (None where not wanted), all fp32."""
    _need(dy, F32, "dy"); _need(x, F32, "x"); _need(w, F32, "w")
    B, K = x.shape
    N = w.shape[0]
    dev = x.device
    dx = torch.empty((B, K), device=dev, dtype=F32) if want_dx else None
    dw = torch.empty((N, K), device=dev, dtype=F32) if want_dw else None
    db = torch.empty((N,), device=dev, dtype=F32) if want_db else None
    with on_device(dev):
        check(lib().b200dt_skinny_linear_bwd(ptr(dy), ptr(x), ptr(w), ptr(dx), ptr(dw), ptr(db), B, K, N,
                                             stream_ptr(dev)), "skinny_linear_bwd")
    return dx, dw, db


def conv2d_fprop(x: torch.Tensor, w: torch.Tensor, stride: int = 1, relu: bool = False,
                 bn_sums: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need(x, BF16, "x"); _need(w, BF16, "w")
    NB, H, W, Cin = x.shape
    Cout, R, S, Cin2 = w.shape
    assert R == S and Cin == Cin2
    y = torch.empty((NB, H // stride, W // stride, Cout), device=x.device, dtype=BF16)
    with on_device(x.device):
        check(lib().b200dt_conv2d_fprop_bf16(ptr(x), ptr(w), ptr(y), NB, H, W, Cin, Cout, R, stride,
                                             int(relu), ptr(bn_sums), stream_ptr(x.device)), "conv2d_fprop")
    return y


def conv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, R: int, stride: int = 1,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need(x, BF16, "x"); _need(dy, BF16, "dy")
    NB, H, W, Cin = x.shape
    Cout = dy.shape[-1]
    if out is None:
        out = torch.empty((Cout, R, R, Cin), device=x.device, dtype=F32)
    with on_device(x.device):
        check(lib().b200dt_conv2d_wgrad_bf16(ptr(x), ptr(dy), ptr(out), NB, H, W, Cin, Cout, R, stride,
                                             stream_ptr(x.device)), "conv2d_wgrad")
    return out


def conv2d_dgrad(dy: torch.Tensor, w: torch.Tensor, x_shape, R: int, stride: int = 1,
                 accumulate_into: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Data gradient; ``w`` is the forward filter ``[Cout, R, R, Cin]`` (no transposed copy needed).
    ``accumulate_into``: a bf16 gradient of the same tensor that already exists (residual branch,
    second consumer); the kernel adds to it IN PLACE in its epilogue and returns it."""
    _need(dy, BF16, "dy"); _need(w, BF16, "w")
    NB, H, W, Cin = x_shape
    Cout = dy.shape[-1]
    if accumulate_into is not None:
        acc = accumulate_into
        _need(acc, BF16, "accumulate_into")
        if tuple(acc.shape) != (NB, H, W, Cin):
            raise ValueError("accumulate_into has the wrong shape")
        with on_device(dy.device):
            rc = lib().b200dt_conv2d_dgrad_acc_bf16(ptr(dy), ptr(w), ptr(acc), NB, H, W, Cin, Cout, R, stride,
                                                    stream_ptr(dy.device))
        if rc == 0:
            return acc
        if rc != -4:
            check(rc, "conv2d_dgrad_acc")
        return acc.add_(conv2d_dgrad(dy, w, x_shape, R, stride))     # output not expressible as a tensor map
    dx = torch.empty((NB, H, W, Cin), device=dy.device, dtype=BF16)
    with on_device(dy.device):
        check(lib().b200dt_conv2d_dgrad_bf16(ptr(dy), ptr(w), ptr(dx), NB, H, W, Cin, Cout, R, stride,
                                             stream_ptr(dy.device)), "conv2d_dgrad")
    return dx


# --------------------------------------------------------------------------- layout
STEM_LDK = 160


def stem_im2col(x: torch.Tensor) -> torch.Tensor:
    """[N,3,H,W] fp32 NCHW -> [N*H/2*W/2, 160] bf16 patches of the 7x7 stride-2 stem."""
    _need(x, F32, "x")
    NB, C, H, W = x.shape
    assert C == 3
    cols = torch.empty((NB * (H // 2) * (W // 2), STEM_LDK), device=x.device, dtype=BF16)
    with on_device(x.device):
        check(lib().b200dt_stem_im2col(ptr(x), ptr(cols), NB, H, W, STEM_LDK, stream_ptr(x.device)),
              "stem_im2col")
    return cols


def stem_supported(x_shape) -> bool:
    """Shapes the implicit-GEMM stem takes: even H, W with (W/2) % 16 == 0 and (H/2) % 8 == 0 (224 and 128 do)."""
    NB, C, H, W = x_shape
    return C == 3 and H % 2 == 0 and W % 2 == 0 and (W // 2) % 16 == 0 and (H // 2) % 8 == 0


def stem_conv_fprop(x: torch.Tensor, w: torch.Tensor, bn_sums: Optional[torch.Tensor] = None) -> torch.Tensor:
    """7x7 s2 p3 stem: x [N,3,H,W] fp32 NCHW, w [64,7,7,3] fp32 -> [N,H/2,W/2,64] bf16 NHWC (no patch matrix)."""
    _need(x, F32, "x"); _need(w, F32, "w")
    NB, _, H, W = x.shape
    assert tuple(w.shape) == (64, 7, 7, 3)
    y = torch.empty((NB, H // 2, W // 2, 64), device=x.device, dtype=BF16)
    with on_device(x.device):
        check(lib().b200dt_stem_conv_fprop(ptr(x), ptr(w), ptr(y), NB, H, W, ptr(bn_sums), stream_ptr(x.device)),
              "stem_conv_fprop")
    return y


def stem_conv_wgrad(x: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    _need(x, F32, "x"); _need(dy, BF16, "dy")
    NB, _, H, W = x.shape
    dw = torch.empty((64, 7, 7, 3), device=x.device, dtype=F32)
    scratch = torch.empty(int(lib().b200dt_stem_wgrad_scratch_floats()), device=x.device, dtype=F32)
    with on_device(x.device):
        check(lib().b200dt_stem_conv_wgrad(ptr(x), ptr(dy), ptr(dw), ptr(scratch), NB, H, W, stream_ptr(x.device)),
              "stem_conv_wgrad")
    return dw


def cast_bf16(src: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need(src, F32, "src")
    if out is None:
        out = torch.empty(src.shape, device=src.device, dtype=BF16)
    with on_device(src.device):
        check(lib().b200dt_cast_f32_to_bf16(ptr(src), ptr(out), src.numel(), stream_ptr(src.device)), "cast")
    return out


def cast_f32(src: torch.Tensor) -> torch.Tensor:
    """bf16 (or fp32) -> fp32 copy on the library's own cast kernel (no ATen elementwise launch)."""
    if src.dtype == F32:
        return src
    _need(src, BF16, "src")
    out = torch.empty(src.shape, device=src.device, dtype=F32)
    with on_device(src.device):
        check(lib().b200dt_cast_to_f32(ptr(src), ptr(out), src.numel(), 1, stream_ptr(src.device)), "cast_f32")
    return out


def transpose_bf16(src: torch.Tensor) -> torch.Tensor:
    _need(src, BF16, "src")
    r, c = src.shape
    out = torch.empty((c, r), device=src.device, dtype=BF16)
    with on_device(src.device):
        check(lib().b200dt_transpose_bf16(ptr(src), ptr(out), r, c, stream_ptr(src.device)), "transpose")
    return out


def nchw_to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    _need(x, F32, "x")
    NB, Cc, H, W = x.shape
    y = torch.empty((NB, H, W, Cc), device=x.device, dtype=BF16)
    with on_device(x.device):
        check(lib().b200dt_nchw_f32_to_nhwc_bf16(ptr(x), ptr(y), NB, Cc, H, W, stream_ptr(x.device)), "nchw_to_nhwc")
    return y


# --------------------------------------------------------------------------- batch norm
def bn_stats(x2d: torch.Tensor) -> torch.Tensor:
    rows, Cc = x2d.shape
    sums = bn_sums_buffer(Cc, x2d.device)
    with on_device(x2d.device):
        check(lib().b200dt_bn_stats(ptr(x2d), ptr(sums), rows, Cc, stream_ptr(x2d.device)), "bn_stats")
    return sums


def bn_apply(x2d, residual, sums, gamma, beta, running_mean, running_var, eps, momentum, relu,
             batches_tracked: Optional[torch.Tensor] = None, relu_mask: Optional[torch.Tensor] = None
             ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``batches_tracked``: optional int64 device scalar the kernel increments (no launch of its own).
    ``relu_mask``: optional uint8 ``[rows * C / 8]`` that receives one bit per output element (> 0);
    ``bn_backward(..., relu_mask=)`` then masks the gradient without re-reading the output."""
    if batches_tracked is not None:
        _need(batches_tracked, torch.int64, "batches_tracked")
    if relu_mask is not None:
        _need(relu_mask, torch.uint8, "relu_mask")
        if relu_mask.numel() != x2d.numel() // 8:
            raise ValueError("relu_mask needs rows * C / 8 bytes")
    rows, Cc = x2d.shape
    y = torch.empty_like(x2d)
    save_mean = torch.empty(Cc, device=x2d.device, dtype=F32)
    save_invstd = torch.empty(Cc, device=x2d.device, dtype=F32)
    with on_device(x2d.device):
        check(lib().b200dt_bn_apply(ptr(x2d), ptr(residual), ptr(y), ptr(sums), ptr(gamma), ptr(beta),
                                    ptr(running_mean), ptr(running_var), ptr(save_mean), ptr(save_invstd),
                                    rows, Cc, eps, momentum, int(relu), ptr(batches_tracked),
                                    ptr(relu_mask), stream_ptr(x2d.device)), "bn_apply")
    return y, save_mean, save_invstd


def bn_apply_eval(x2d, residual, gamma, beta, mean, var, eps, relu) -> torch.Tensor:
    rows, Cc = x2d.shape
    y = torch.empty_like(x2d)
    with on_device(x2d.device):
        check(lib().b200dt_bn_apply_eval(ptr(x2d), ptr(residual), ptr(y), ptr(gamma), ptr(beta), ptr(mean),
                                         ptr(var), rows, Cc, eps, int(relu), stream_ptr(x2d.device)), "bn_eval")
    return y


def bn_backward(dy2d, y2d, x2d, save_mean, save_invstd, gamma, relu: bool, want_dres: bool, beta=None,
                relu_mask: Optional[torch.Tensor] = None):
    """Returns (dx, dgamma, dbeta, dres or None). With ``relu`` and ``y2d is None`` the ReLU mask is
    recomputed from ``x2d`` (needs ``beta``; only valid when the forward had no residual input);
    with ``relu_mask`` (from ``bn_apply``) it is read from one byte per eight channels."""
    rows, Cc = x2d.shape
    dev = x2d.device
    if relu and relu_mask is not None:
        dres = torch.empty_like(dy2d)
        dx = torch.empty_like(x2d)
        sums = bn_sums_buffer(Cc, dev)
        dgamma = torch.empty(Cc, device=dev, dtype=F32)
        dbeta = torch.empty(Cc, device=dev, dtype=F32)
        with on_device(dev):
            st = stream_ptr(dev)
            check(lib().b200dt_bn_bwd_reduce(ptr(dy2d), ptr(relu_mask), ptr(x2d), ptr(save_mean), ptr(save_invstd),
                                             ptr(gamma), ptr(beta), ptr(sums), ptr(dres), rows, Cc, 2, st),
                  "bn_bwd_reduce")
            check(lib().b200dt_bn_bwd_apply(ptr(dres), None, ptr(x2d), ptr(save_mean), ptr(save_invstd),
                                            ptr(gamma), ptr(beta), ptr(sums), ptr(dx), ptr(dgamma), ptr(dbeta),
                                            rows, Cc, 0, st), "bn_bwd_apply")
        return dx, dgamma, dbeta, (dres if want_dres else None)
    sums = bn_sums_buffer(Cc, dev)
    dres = torch.empty_like(dy2d) if want_dres else None
    dx = torch.empty_like(x2d)
    dgamma = torch.empty(Cc, device=dev, dtype=F32)
    dbeta = torch.empty(Cc, device=dev, dtype=F32)
    with on_device(dev):
        st = stream_ptr(dev)
        check(lib().b200dt_bn_bwd_reduce(ptr(dy2d), ptr(y2d), ptr(x2d), ptr(save_mean), ptr(save_invstd),
                                         ptr(gamma), ptr(beta), ptr(sums), ptr(dres), rows, Cc, int(relu), st),
              "bn_bwd_reduce")
        if want_dres:   # masked gradient already materialised: reuse it, skip the y read
            check(lib().b200dt_bn_bwd_apply(ptr(dres), None, ptr(x2d), ptr(save_mean), ptr(save_invstd),
                                            ptr(gamma), ptr(beta), ptr(sums), ptr(dx), ptr(dgamma), ptr(dbeta),
                                            rows, Cc, 0, st), "bn_bwd_apply")
        else:
            check(lib().b200dt_bn_bwd_apply(ptr(dy2d), ptr(y2d), ptr(x2d), ptr(save_mean), ptr(save_invstd),
                                            ptr(gamma), ptr(beta), ptr(sums), ptr(dx), ptr(dgamma), ptr(dbeta),
                                            rows, Cc, int(relu), st), "bn_bwd_apply")
    return dx, dgamma, dbeta, dres


# --------------------------------------------------------------------------- pooling
def maxpool_fwd(x: torch.Tensor):
    NB, H, W, Cc = x.shape
    y = torch.empty((NB, H // 2, W // 2, Cc), device=x.device, dtype=BF16)
    arg = torch.empty((NB, H // 2, W // 2, Cc), device=x.device, dtype=torch.uint8)
    with on_device(x.device):
        check(lib().b200dt_maxpool3x3s2_fwd(ptr(x), ptr(y), ptr(arg), NB, H, W, Cc, stream_ptr(x.device)), "maxpool")
    return y, arg


def maxpool_bwd(dy: torch.Tensor, arg: torch.Tensor, x_shape) -> torch.Tensor:
    NB, H, W, Cc = x_shape
    dx = torch.empty(x_shape, device=dy.device, dtype=BF16)
    with on_device(dy.device):
        check(lib().b200dt_maxpool3x3s2_bwd(ptr(dy), ptr(arg), ptr(dx), NB, H, W, Cc, stream_ptr(dy.device)), "maxpool_bwd")
    return dx


def avgpool_fwd(x: torch.Tensor) -> torch.Tensor:
    NB, H, W, Cc = x.shape
    y = torch.empty((NB, Cc), device=x.device, dtype=BF16)
    with on_device(x.device):
        check(lib().b200dt_avgpool_fwd(ptr(x), ptr(y), NB, H * W, Cc, stream_ptr(x.device)), "avgpool")
    return y


def avgpool_bwd(dy: torch.Tensor, x_shape) -> torch.Tensor:
    NB, H, W, Cc = x_shape
    dx = torch.empty(x_shape, device=dy.device, dtype=BF16)
    with on_device(dy.device):
        check(lib().b200dt_avgpool_bwd(ptr(dy), ptr(dx), NB, H * W, Cc, stream_ptr(dy.device)), "avgpool_bwd")
    return dx


# --------------------------------------------------------------------------- losses
_scratch = {}


def loss_scratch(device) -> torch.Tensor:
    """Per-(device, stream) zero-initialised scratch for the two-level loss reductions."""
    key = (torch.device(device).index, torch.cuda.current_stream(device).cuda_stream)
    s = _scratch.get(key)
    if s is None:
        n = int(lib().b200dt_loss_scratch_bytes())
        s = torch.zeros(n, device=device, dtype=torch.uint8)
        _scratch[key] = s
    return s


def _loss_common(kind: str, pred: torch.Tensor, other, grad_scale: float, want_grad: bool):
    dev = pred.device
    is_bf16 = pred.dtype == BF16
    if pred.dtype not in (BF16, F32) or not pred.is_contiguous():
        raise ValueError("loss input must be contiguous fp32 or bf16")
    loss = torch.empty((), device=dev, dtype=F32)
    grad = torch.empty_like(pred) if want_grad else None
    with on_device(dev):
        sc = loss_scratch(dev)
        st = stream_ptr(dev)
        L = lib()
        if kind == "mse":
            other = other.to(dtype=pred.dtype).contiguous()
            check(L.b200dt_mse_loss(ptr(pred), ptr(other), ptr(grad), ptr(loss), pred.numel(), grad_scale,
                                    int(is_bf16), ptr(sc), st), "mse")
        elif kind == "sum":
            check(L.b200dt_sum_loss(ptr(pred), ptr(grad), ptr(loss), pred.numel(), grad_scale, int(is_bf16),
                                    ptr(sc), st), "sum")
        elif kind == "ce_prob":
            B, Cc = pred.shape
            other = other.to(dtype=F32).contiguous()
            check(L.b200dt_ce_prob_loss(ptr(pred), ptr(other), ptr(grad), ptr(loss), B, Cc, grad_scale,
                                        int(is_bf16), ptr(sc), st), "ce_prob")
        elif kind == "ce_index":
            B, Cc = pred.shape
            other = other.to(dtype=torch.int64).contiguous()
            check(L.b200dt_ce_index_loss(ptr(pred), ptr(other), ptr(grad), ptr(loss), B, Cc, grad_scale,
                                         int(is_bf16), ptr(sc), st), "ce_index")
        else:  # pragma: no cover
            raise ValueError(kind)
    return loss, grad


def mse_loss_fwd_bwd(pred, target, grad_scale=1.0, want_grad=True):
    return _loss_common("mse", pred, target, grad_scale, want_grad)


def sum_loss_fwd_bwd(x, grad_scale=1.0, want_grad=True):
    return _loss_common("sum", x, None, grad_scale, want_grad)


def ce_prob_fwd_bwd(logits, probs, grad_scale=1.0, want_grad=True):
    return _loss_common("ce_prob", logits, probs, grad_scale, want_grad)


def ce_index_fwd_bwd(logits, labels, grad_scale=1.0, want_grad=True):
    return _loss_common("ce_index", logits, labels, grad_scale, want_grad)
