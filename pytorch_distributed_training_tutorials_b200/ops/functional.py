"""Differentiable ops on top of the sm_100a kernels (``torch.autograd.Function`` per op).

Conventions: activations are NHWC bf16 tensors ``[N, H, W, C]``; parameters are fp32 masters
(what the optimizer and DDP see) and are cast to bf16 operands on the fly; weight gradients come
back in fp32 straight from the tensor-core accumulators.
"""
from __future__ import annotations

from typing import Optional

import os

import torch

from . import native as NV

# "implicit": the stem runs as an implicit GEMM (csrc/stem.cu, no patch matrix in memory); "im2col": explicit gather
# + GEMM.  The default follows the measurements in profiles/ (B200DT_STEM overrides it).
STEM_MODE = os.environ.get("B200DT_STEM", "implicit")

BF16 = torch.bfloat16
F32 = torch.float32


def _shadow(w: torch.Tensor) -> torch.Tensor:
    """bf16 operand copy of an fp32 master weight, cached until the master changes."""
    cache = getattr(w, "_b200_shadow", None)
    ver = w._version
    if cache is not None and cache[0] == ver and cache[1].device == w.device:
        return cache[1]
    s = NV.cast_bf16(w.detach().contiguous())
    try:
        w._b200_shadow = (ver, s)
    except Exception:  # pragma: no cover - tensors that refuse attributes
        pass
    return s


def _grad_out(p: torch.Tensor) -> Optional[torch.Tensor]:
    """Destination for a freshly computed weight gradient when DDP keeps gradients as views of its
    symmetric arena (``gradient_as_bucket_view``): the kernel writes straight into the bucket slot and
    autograd adopts that tensor as ``.grad``, so the all-reduce needs no copy-in.  None otherwise, and
    while a gradient is being accumulated (then autograd must add, not overwrite)."""
    slot = getattr(p, "_b200_grad_slot", None)
    if slot is None or p.grad is not None:
        return None
    region, off = slot
    if region.dtype != F32 or region.device != p.device:
        return None
    return region[off:off + p.numel()].view(p.shape)


class _Fn(torch.autograd.Function):
    """``Function.apply`` minus the functorch-wrapper bookkeeping torch runs on every call (3 us per op, ~100 ops per
    ResNet-50 step: eager steps here are bound by the host).  vmap / functorch transforms over these ops are not
    supported either way: they are opaque kernel launches."""

    @classmethod
    def apply(cls, *args):
        return super(torch.autograd.Function, cls).apply(*args)


class _Linear(_Fn):
    @staticmethod
    def forward(ctx, x, weight, bias, relu, out_fp32):
        wb = _shadow(weight)
        x2 = x.reshape(-1, x.shape[-1])
        if x2.dtype != BF16:
            x2 = x2.to(BF16)
        x2 = x2.contiguous()
        y = NV.gemm_tn(x2, wb, bias=bias.detach() if bias is not None else None, relu=relu,
                       out_dtype=F32 if out_fp32 else BF16)
        ctx.save_for_backward(x2, weight, y if relu else None)
        ctx.has_bias = bias is not None
        ctx.relu = relu
        ctx.x_shape = x.shape
        ctx.x_dtype = x.dtype
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, y = ctx.saved_tensors
        N, Kf = weight.shape
        dy2 = dy.reshape(-1, N)
        dx = dw = db = None
        want_db = ctx.has_bias and ctx.needs_input_grad[2]
        if N % 8:
            # an output width the TMA row pitch cannot take (e.g. a 1001-way head): the backward runs on the
            # fp32 SIMT kernels of csrc/skinny.cu -- still this library's code, no cuBLAS call
            if ctx.relu:
                dy2 = dy2 * (y > 0)
            dyf = dy2.contiguous() if dy2.dtype == F32 else NV.cast_f32(dy2.contiguous())
            dx, dw, db = NV.skinny_linear_bwd(dyf, NV.cast_f32(x2), weight.detach().contiguous(),
                                              ctx.needs_input_grad[0], ctx.needs_input_grad[1], want_db)
            if dx is not None:
                dx = dx.view(ctx.x_shape).to(ctx.x_dtype)
            return dx, dw, db, None, None
        if (ctx.relu or want_db) and (not ctx.relu or y.dtype == BF16):
            # ReLU mask and bias gradient in one pass (bf16 in, bf16 out, fp32 column sums)
            dyb, db = NV.relu_bwd_bias(dy2.to(BF16).contiguous(), y.reshape(-1, N) if ctx.relu else None, want_db)
            dy2 = dyb
            want_db = False
        else:
            if ctx.relu:
                dy2 = dy2 * (y > 0)
            dyb = dy2.to(BF16).contiguous()
        if ctx.needs_input_grad[0]:
            # W [out, in] is consumed in place as the MN-major B operand: no transposed copy
            dx = NV.gemm_nn(dyb, _shadow(weight)).view(ctx.x_shape).to(ctx.x_dtype)
        if ctx.needs_input_grad[1]:
            dw = NV.gemm_nt_reduce(dyb, x2, out=_grad_out(weight))
        if want_db:
            db = dy2.float().sum(0)
        return dx, dw, db, None, None


class _SkinnyLinear(_Fn):
    """fp32 Linear for feature counts the tensor-core path cannot take (the tutorial's Linear(20,1),
    (32,2), (10000,10), (10,5); ddp_gpus.py:81, 01.data_parallel.ipynb:171, 03.model_parallel.ipynb:444-446)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1]).float().contiguous()
        w = weight.detach().float().contiguous()
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        y = NV.skinny_linear_fwd(x2, w, bias.detach().float().contiguous() if bias is not None else None)
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, weight.shape[0]).float().contiguous()
        dx, dw, db = NV.skinny_linear_bwd(dy2, x2, weight.detach().float().contiguous(), ctx.needs_input_grad[0],
                                          ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2])
        if dx is not None:
            dx = dx.view(ctx.x_shape).to(ctx.x_dtype)
        return dx, dw, db


def skinny_linear(x, weight, bias=None):
    """fp32 ``x @ weight^T + bias`` for any (in_features, out_features) on the SIMT kernel."""
    return _SkinnyLinear.apply(x, weight, bias)


_cpp_linear = None          # the C++ extension once its kernel addresses are registered; False = not available / off


def _native_linear():
    """csrc/reducer/linear.cpp: the same forward / backward as ``_Linear`` as a C++ autograd node, so the backward
    passes of DataParallel replicas (one autograd thread per device) do not serialise on the GIL.
    ``B200DT_CPP_LINEAR=0`` keeps the Python node."""
    global _cpp_linear
    if _cpp_linear is None:
        _cpp_linear = False
        if os.environ.get("B200DT_CPP_LINEAR", "1") not in ("0", ""):
            try:
                import ctypes as C
                from .._lib import lib
                from ..parallel._native_reducer import load
                mod = load()
                h = lib()._handle
                addr = lambda name: C.cast(getattr(h, name), C.c_void_p).value
                mod.linear_set_fns(addr("b200dt_gemm_tn_bf16"), addr("b200dt_gemm_nn_bf16"),
                                   addr("b200dt_gemm_nt_reduce_bf16"), addr("b200dt_relu_bwd_bias"))
                _cpp_linear = mod
            except Exception:      # noqa: BLE001 -- no extension / old build: the Python node does the same work
                _cpp_linear = False
    return _cpp_linear


def linear(x, weight, bias=None, relu: bool = False, out_fp32: bool = False):
    """y = x @ weight^T + bias on tcgen05; needs in_features % 8 == 0 (TMA row pitch)."""
    if (x.is_cuda and weight.shape[0] % 8 == 0 and weight.shape[1] % 8 == 0 and weight.dtype == F32
            and getattr(weight, "_b200_grad_slot", None) is None and (bias is None or bias.dtype == F32)):
        mod = _native_linear()
        if mod:
            from .. import _lib
            _lib.launch_count += 1 + (3 if torch.is_grad_enabled() and (weight.requires_grad or x.requires_grad) else 0)
            return mod.linear(x, weight, _shadow(weight), bias, bool(relu), bool(out_fp32))
    return _Linear.apply(x, weight, bias, relu, out_fp32)


class _Conv2d(_Fn):
    @staticmethod
    def forward(ctx, x, weight, stride, with_stats):
        wb = _shadow(weight)
        ctx.save_for_backward(x, weight)
        ctx.stride = stride
        ctx.with_stats = with_stats
        ctx.set_materialize_grads(False)      # no zero tensor for the statistics output in backward
        if with_stats:
            # per-channel sum / sum of squares of the output, accumulated in the conv epilogue
            sums = NV.bn_sums_buffer(weight.shape[0], x.device)
            y = NV.conv2d_fprop(x, wb, stride=stride, bn_sums=sums)
            ctx.mark_non_differentiable(sums)
            return y, sums
        return NV.conv2d_fprop(x, wb, stride=stride)

    @staticmethod
    def backward(ctx, dy, *unused):
        if dy is None:
            return None, None, None, None
        x, weight = ctx.saved_tensors
        R = weight.shape[1]
        dy = dy.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = NV.conv2d_dgrad(dy, _shadow(weight), x.shape, R, stride=ctx.stride)
        if ctx.needs_input_grad[1]:
            dw = NV.conv2d_wgrad(x, dy, R, stride=ctx.stride, out=_grad_out(weight))
        return dx, dw, None, None


def conv2d(x, weight, stride: int = 1, with_stats: bool = False):
    """NHWC conv, weight ``[Cout, R, R, Cin]``, R in {1,3}, pad R//2, stride in {1,2}.
    ``with_stats=True`` returns ``(y, sums)``: the BatchNorm statistics of ``y`` come out of the
    conv epilogue for free; hand ``sums`` to ``batch_norm(..., sums=sums)``."""
    return _Conv2d.apply(x, weight, stride, with_stats)


class _ConvSkip(_Fn):
    """``(conv(x, w), x)``: the convolution that opens a residual block together with the block's
    identity branch, so that backward can add the two gradients of ``x`` inside the dgrad epilogue
    (TMA reduce-add into the incoming skip gradient) instead of leaving a separate add to autograd."""

    @staticmethod
    def forward(ctx, x, weight, stride, with_stats):
        wb = _shadow(weight)
        ctx.save_for_backward(x, weight)
        ctx.stride = stride
        ctx.set_materialize_grads(False)
        if with_stats:
            sums = NV.bn_sums_buffer(weight.shape[0], x.device)
            y = NV.conv2d_fprop(x, wb, stride=stride, bn_sums=sums)
            ctx.mark_non_differentiable(sums)
            return y, sums, x
        return NV.conv2d_fprop(x, wb, stride=stride), None, x

    @staticmethod
    def backward(ctx, dy, _dsums, dskip):
        x, weight = ctx.saved_tensors
        R = weight.shape[1]
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if dy is None:
                dx = dskip
            else:
                acc = dskip.contiguous() if dskip is not None else None
                dx = NV.conv2d_dgrad(dy.contiguous(), _shadow(weight), x.shape, R, stride=ctx.stride,
                                     accumulate_into=acc)
        if ctx.needs_input_grad[1] and dy is not None:
            dw = NV.conv2d_wgrad(x, dy.contiguous(), R, stride=ctx.stride, out=_grad_out(weight))
        return dx, dw, None, None


def conv2d_skip(x, weight, stride: int = 1, with_stats: bool = False):
    """Returns ``(y, sums or None, skip)`` with ``skip`` an alias of ``x`` for the residual branch.
    The gradient that comes back through ``skip`` is consumed IN PLACE (it must not be shared; the
    fused BatchNorm backward hands out a fresh tensor, which is what the models here rely on)."""
    return _ConvSkip.apply(x, weight, stride, with_stats)


class _ConvDual(_Fn):
    """Two convolutions of the same input (a block's first conv and its downsample projection):
    backward writes the first data gradient and adds the second one in its epilogue."""

    @staticmethod
    def forward(ctx, x, w1, w2, stride1, stride2):
        ctx.save_for_backward(x, w1, w2)
        ctx.strides = (stride1, stride2)
        ctx.set_materialize_grads(False)
        s1 = NV.bn_sums_buffer(w1.shape[0], x.device)
        s2 = NV.bn_sums_buffer(w2.shape[0], x.device)
        y1 = NV.conv2d_fprop(x, _shadow(w1), stride=stride1, bn_sums=s1)
        y2 = NV.conv2d_fprop(x, _shadow(w2), stride=stride2, bn_sums=s2)
        ctx.mark_non_differentiable(s1, s2)
        return y1, s1, y2, s2

    @staticmethod
    def backward(ctx, dy1, _d1, dy2, _d2):
        x, w1, w2 = ctx.saved_tensors
        st1, st2 = ctx.strides
        dy1 = dy1.contiguous() if dy1 is not None else None
        dy2 = dy2.contiguous() if dy2 is not None else None
        dx = dw1 = dw2 = None
        if ctx.needs_input_grad[0]:
            if dy1 is not None:
                dx = NV.conv2d_dgrad(dy1, _shadow(w1), x.shape, w1.shape[1], stride=st1)
            if dy2 is not None:
                dx = NV.conv2d_dgrad(dy2, _shadow(w2), x.shape, w2.shape[1], stride=st2, accumulate_into=dx)
        if ctx.needs_input_grad[1] and dy1 is not None:
            dw1 = NV.conv2d_wgrad(x, dy1, w1.shape[1], stride=st1, out=_grad_out(w1))
        if ctx.needs_input_grad[2] and dy2 is not None:
            dw2 = NV.conv2d_wgrad(x, dy2, w2.shape[1], stride=st2, out=_grad_out(w2))
        return dx, dw1, dw2, None, None


def conv2d_dual(x, w1, w2, stride1: int = 1, stride2: int = 1):
    """``(conv(x, w1), sums1, conv(x, w2), sums2)`` with one fused data gradient for ``x``."""
    return _ConvDual.apply(x, w1, w2, stride1, stride2)


def _stem_operand(w: torch.Tensor) -> torch.Tensor:
    """The stem filter as the GEMM's B operand: bf16 ``[Cout, 160]`` (K = 147 zero-padded to the TMA row pitch),
    cached until the fp32 master changes -- it used to be rebuilt with a fill and a strided copy every forward."""
    cache = getattr(w, "_b200_stem_operand", None)
    if cache is not None and cache[0] == w._version and cache[1].device == w.device:
        return cache[1]
    wk = torch.zeros((w.shape[0], NV.STEM_LDK), device=w.device, dtype=BF16)
    wk[:, :147] = w.detach().reshape(w.shape[0], 147)
    try:
        w._b200_stem_operand = (w._version, wk)
    except Exception:  # pragma: no cover
        pass
    return wk


class _StemConv(_Fn):
    """7x7 stride-2 pad-3 convolution of the fp32 NCHW network input (Cin = 3)."""

    @staticmethod
    def forward(ctx, x, weight, with_stats):
        NB, _, H, W = x.shape
        ctx.implicit = (STEM_MODE == "implicit" and weight.shape[0] == 64 and NV.stem_supported(x.shape)
                        and x.dtype == F32)
        if ctx.implicit:
            # implicit GEMM: the patch matrix exists only tile by tile in shared memory (csrc/stem.cu)
            xc = x.contiguous()
            sums = NV.bn_sums_buffer(64, x.device) if with_stats else None
            ctx.set_materialize_grads(False)
            y = NV.stem_conv_fprop(xc, weight.detach().contiguous(), bn_sums=sums)
            ctx.save_for_backward(xc, weight)
            if with_stats:
                ctx.mark_non_differentiable(sums)
                return y, sums
            return y
        cols = NV.stem_im2col(x.contiguous())
        wk = _stem_operand(weight)
        sums = NV.bn_sums_buffer(weight.shape[0], x.device) if with_stats else None
        ctx.set_materialize_grads(False)
        y = NV.gemm_tn(cols, wk, bn_sums=sums)
        # the patch matrix is kept for the weight gradient (memory is cheap on a 180 GB part;
        # re-gathering it costs a full extra pass over the input)
        ctx.save_for_backward(cols, weight)
        y = y.view(NB, H // 2, W // 2, weight.shape[0])
        if with_stats:
            ctx.mark_non_differentiable(sums)
            return y, sums
        return y

    @staticmethod
    def backward(ctx, dy, *unused):
        cols, weight = ctx.saved_tensors
        dw = None
        if ctx.implicit:
            if ctx.needs_input_grad[1] and dy is not None:
                dw = NV.stem_conv_wgrad(cols, dy.contiguous())          # `cols` is the network input here
            return None, dw, None
        if ctx.needs_input_grad[1] and dy is not None:
            dy2 = dy.reshape(-1, weight.shape[0]).contiguous()
            dwk = NV.gemm_nt_reduce(dy2, cols)               # [64, 160]
            dw = dwk[:, :147].reshape(weight.shape).contiguous()
        return None, dw, None


def stem_conv(x_nchw_f32, weight, with_stats: bool = False):
    """weight ``[64, 7, 7, 3]`` (OHWI)."""
    return _StemConv.apply(x_nchw_f32, weight, with_stats)


class _BatchNorm(_Fn):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu, training, sums,
                tracked=None):
        shp = x.shape
        Cc = shp[-1]
        x2 = x.reshape(-1, Cc)
        r2 = residual.reshape(-1, Cc) if residual is not None else None
        mask = None
        if training:
            if sums is None:
                sums = NV.bn_stats(x2)
            if relu and residual is not None and x2.numel() % 8 == 0:
                # with a residual input the ReLU mask cannot be recomputed from x: keep one BIT per element
                mask = torch.empty(x2.numel() // 8, device=x2.device, dtype=torch.uint8)
            y, mean, invstd = NV.bn_apply(x2, r2, sums, gamma.detach(), beta.detach(), running_mean,
                                          running_var, eps, momentum, relu, tracked, mask)
        else:
            y = NV.bn_apply_eval(x2, r2, gamma.detach(), beta.detach(), running_mean, running_var, eps, relu)
            mean = running_mean
            invstd = torch.rsqrt(running_var + eps)
        # without a residual input the ReLU mask is recomputed from x in backward, with one the bit
        # mask is kept: the output y itself is never saved by this op
        keep_y = relu and residual is not None and mask is None
        ctx.save_for_backward(x2, y if keep_y else None, gamma, beta, mean, invstd, mask)
        ctx.relu = relu
        ctx.has_res = residual is not None
        ctx.shape = shp
        ctx.training = training
        return y.view(shp)

    @staticmethod
    def backward(ctx, dy):
        x2, y, gamma, beta, mean, invstd, mask = ctx.saved_tensors
        Cc = ctx.shape[-1]
        dy2 = dy.reshape(-1, Cc).contiguous()
        if not ctx.training:
            raise RuntimeError("backward through eval-mode BatchNorm is not supported")
        want_res = ctx.has_res and ctx.needs_input_grad[3]
        dx, dgamma, dbeta, dres = NV.bn_backward(dy2, y, x2, mean, invstd, gamma.detach(), ctx.relu, want_res,
                                                 beta=beta.detach(), relu_mask=mask)
        return (dx.view(ctx.shape), dgamma, dbeta, dres.view(ctx.shape) if dres is not None else None,
                None, None, None, None, None, None, None, None)


def batch_norm(x, gamma, beta, running_mean, running_var, residual=None, eps: float = 1e-5,
               momentum: float = 0.1, relu: bool = False, training: bool = True, sums=None,
               num_batches_tracked=None):
    """y = act(BN(x) + residual) with batch statistics; updates running stats in place.
    ``sums``: statistics already produced by the conv that wrote ``x`` (skips the stats pass).
    ``num_batches_tracked``: int64 device scalar incremented inside the apply kernel when training."""
    return _BatchNorm.apply(x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu, training,
                            sums, num_batches_tracked)


class _MaxPool(_Fn):
    @staticmethod
    def forward(ctx, x):
        y, arg = NV.maxpool_fwd(x.contiguous())
        ctx.save_for_backward(arg)
        ctx.shape = x.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        return NV.maxpool_bwd(dy.contiguous(), arg, ctx.shape)


def max_pool_3x3_s2(x):
    return _MaxPool.apply(x)


class _AvgPool(_Fn):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = x.shape
        return NV.avgpool_fwd(x.contiguous())

    @staticmethod
    def backward(ctx, dy):
        return NV.avgpool_bwd(dy.contiguous(), ctx.shape)


def global_avg_pool(x):
    """[N,H,W,C] -> [N,C]."""
    return _AvgPool.apply(x)


class _Loss(_Fn):
    @staticmethod
    def forward(ctx, kind, pred, target):
        pred_c = pred.contiguous()
        want = pred.requires_grad
        if kind == "mse":
            loss, grad = NV.mse_loss_fwd_bwd(pred_c, target, 1.0, want)
        elif kind == "sum":
            loss, grad = NV.sum_loss_fwd_bwd(pred_c, 1.0, want)
        elif kind == "ce_prob":
            loss, grad = NV.ce_prob_fwd_bwd(pred_c, target, 1.0, want)
        else:
            loss, grad = NV.ce_index_fwd_bwd(pred_c, target, 1.0, want)
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return None, (grad * g.to(grad.dtype)) if grad is not None else None, None


def mse_loss(pred, target):
    return _Loss.apply("mse", pred, target)


def sum_loss(x):
    return _Loss.apply("sum", x, None)


def cross_entropy(logits, target):
    """Mean cross entropy; ``target`` is either class indices (int64 ``[B]``) or probabilities
    (float, same shape as ``logits``, the form the reference uses at ddp_gpus.py:37)."""
    if target.dtype in (torch.int64, torch.int32):
        return _Loss.apply("ce_index", logits, target)
    return _Loss.apply("ce_prob", logits, target)
