"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_training_tutorials_b200.ops import native as NV


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a = a.float(); b = b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def nhwc(x_nchw):
    return x_nchw.permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 128, 128), (1000, 1000, 2048), (4096, 256, 512),
                                   (300, 72, 200), (32, 8, 24), (8192, 2048, 1024), (20000, 512, 64)])
def test_gemm_tn(M, N, K):
    torch.manual_seed(0)
    a = torch.randn(M, K, device=dev()).bfloat16()
    b = torch.randn(N, K, device=dev()).bfloat16()
    bias = torch.randn(N, device=dev())
    ref = a.float() @ b.float().t()
    out = NV.gemm_tn(a, b)
    assert rel_err(out, ref) < 1e-2
    out2 = NV.gemm_tn(a, b, bias=bias, relu=True, out_dtype=torch.float32)
    assert rel_err(out2, torch.relu(ref + bias)) < 2e-3


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (2048, 4096, 4096), (300, 72, 200), (16, 2048, 1000), (1000, 1000, 2048)])
def test_gemm_nn(M, N, K):
    torch.manual_seed(0)
    a = torch.randn(M, K, device=dev()).bfloat16()
    b = torch.randn(K, N, device=dev()).bfloat16()
    out = NV.gemm_nn(a, b, out_dtype=torch.float32)
    assert rel_err(out, a.float() @ b.float()) < 2e-3


@pytest.mark.parametrize("R,M,N", [(64, 64, 64), (4096, 128, 128), (1000, 1000, 2048), (777, 72, 200),
                                   (100000, 64, 64), (32768, 256, 64)])
def test_gemm_nt_reduce(R, M, N):
    torch.manual_seed(0)
    a = torch.randn(R, M, device=dev()).bfloat16()
    b = torch.randn(R, N, device=dev()).bfloat16()
    ref = a.float().t() @ b.float()
    out = NV.gemm_nt_reduce(a, b)
    assert rel_err(out, ref) < 2e-3


CONVS = [  # (NB, H, W, Cin, Cout, R, stride)
    (8, 56, 56, 64, 64, 1, 1), (8, 56, 56, 64, 64, 3, 1), (8, 56, 56, 64, 256, 1, 1),
    (8, 56, 56, 256, 128, 1, 1), (8, 56, 56, 128, 128, 3, 2), (8, 56, 56, 256, 512, 1, 2),
    (32, 28, 28, 128, 128, 3, 1), (32, 28, 28, 256, 256, 3, 2), (32, 14, 14, 256, 256, 3, 1),
    (32, 14, 14, 512, 512, 3, 2), (128, 7, 7, 512, 512, 3, 1), (128, 7, 7, 512, 2048, 1, 1),
    (5, 28, 28, 128, 128, 3, 1), (3, 14, 14, 1024, 256, 1, 1), (130, 7, 7, 2048, 512, 1, 1),
]


def conv_ref(x, w, stride):
    # x: NHWC bf16, w: [Cout,R,S,Cin] bf16 -> fp32 NCHW reference
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)
    y = F.conv2d(xr, wr, stride=stride, padding=w.shape[1] // 2)
    return xr, wr, y


@pytest.mark.parametrize("cfg", CONVS)
def test_conv_fprop(cfg):
    NB, H, W, Cin, Cout, R, s = cfg
    torch.manual_seed(1)
    x = torch.randn(NB, H, W, Cin, device=dev()).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=dev()) / (R * Cin ** 0.5)).bfloat16()
    _, _, yref = conv_ref(x, w, s)
    y = NV.conv2d_fprop(x, w, stride=s)
    assert y.shape == (NB, H // s, W // s, Cout)
    assert rel_err(y, nhwc(yref)) < 1e-2


@pytest.mark.parametrize("cfg", CONVS)
def test_conv_dgrad_wgrad(cfg):
    NB, H, W, Cin, Cout, R, s = cfg
    torch.manual_seed(2)
    x = torch.randn(NB, H, W, Cin, device=dev()).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=dev()) / (R * Cin ** 0.5)).bfloat16()
    dy = torch.randn(NB, H // s, W // s, Cout, device=dev()).bfloat16()
    xr, wr, yref = conv_ref(x, w, s)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    dw = NV.conv2d_wgrad(x, dy, R, stride=s)
    assert rel_err(dw, wr.grad.permute(0, 2, 3, 1)) < 5e-3
    dx = NV.conv2d_dgrad(dy, w, x.shape, R, stride=s)
    assert rel_err(dx, nhwc(xr.grad)) < 1e-2


@pytest.mark.parametrize("cfg", [(8, 56, 56, 256, 64, 1, 1), (8, 56, 56, 256, 512, 1, 2), (32, 14, 14, 1024, 256, 1, 1),
                                 (5, 28, 28, 128, 128, 3, 2), (130, 7, 7, 2048, 512, 1, 1), (32, 28, 28, 128, 128, 3, 1)])
def test_conv_dgrad_accumulates_into_existing_gradient(cfg):
    """dx += dgrad(dy): the residual-branch gradient is added inside the epilogue (TMA reduce-add)."""
    NB, H, W, Cin, Cout, R, s = cfg
    torch.manual_seed(12)
    w = (torch.randn(Cout, R, R, Cin, device=dev()) / (R * Cin ** 0.5)).bfloat16()
    dy = torch.randn(NB, H // s, W // s, Cout, device=dev()).bfloat16()
    skip = torch.randn(NB, H, W, Cin, device=dev()).bfloat16()
    plain = NV.conv2d_dgrad(dy, w, (NB, H, W, Cin), R, stride=s)
    want = skip.float() + plain.float()
    acc = skip.clone()
    out = NV.conv2d_dgrad(dy, w, (NB, H, W, Cin), R, stride=s, accumulate_into=acc)
    assert out.data_ptr() == acc.data_ptr()
    assert rel_err(out, want) < 6e-3                 # one bf16 rounding of the sum


def test_conv_skip_and_dual_functions_match_unfused_autograd():
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    torch.manual_seed(13)
    x0 = torch.randn(8, 28, 28, 256, device=dev()).bfloat16()
    w1 = (torch.randn(64, 1, 1, 256, device=dev()) / 16).requires_grad_(True)
    w2 = (torch.randn(512, 1, 1, 256, device=dev()) / 16).requires_grad_(True)
    for fused in (False, True):
        x = x0.clone().requires_grad_(True)
        if fused:
            y, _, skip = Fn.conv2d_skip(x, w1, 1, True)
            a, _, b, _ = Fn.conv2d_dual(x, w1, w2, 1, 2)
        else:
            y, skip = Fn.conv2d(x, w1), x
            a, b = Fn.conv2d(x, w1), Fn.conv2d(x, w2, stride=2)
        loss = (y.float() ** 2).mean() + (skip.float() * 0.37).sum() / skip.numel() + (a.float() * 0.5).mean() \
            + (b.float() ** 2).mean()
        loss.backward()
        if fused:
            got = (x.grad.clone(), w1.grad.clone(), w2.grad.clone())
        else:
            ref = (x.grad.clone(), w1.grad.clone(), w2.grad.clone())
        w1.grad = None; w2.grad = None
    for g, r in zip(got, ref):
        assert rel_err(g, r) < 1e-2


@pytest.mark.parametrize("cfg", [(8, 56, 56, 64, 256, 1, 1), (32, 14, 14, 256, 256, 3, 1), (5, 28, 28, 128, 128, 3, 2), (130, 7, 7, 2048, 512, 1, 1)])
def test_conv_fused_bn_statistics(cfg):
    NB, H, W, Cin, Cout, R, s = cfg
    torch.manual_seed(7)
    x = torch.randn(NB, H, W, Cin, device=dev()).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=dev()) / (R * Cin ** 0.5)).bfloat16()
    sums = NV.bn_sums_buffer(Cout, dev())
    y = NV.conv2d_fprop(x, w, stride=s, bn_sums=sums)
    ref = NV.bn_stats(y.view(-1, Cout))
    yf = y.float().view(-1, Cout)
    assert torch.allclose(sums[:Cout], yf.sum(0), rtol=2e-3, atol=2e-2 * yf.abs().sum(0).max().item() / yf.shape[0] ** 0.5 + 1e-2)
    assert torch.allclose(sums[:2 * Cout], ref[:2 * Cout], rtol=2e-3, atol=1.0)
    assert rel_err(sums[Cout:2 * Cout], (yf * yf).sum(0)) < 2e-3


def test_stem_im2col_conv():
    torch.manual_seed(3)
    x = torch.randn(4, 3, 64, 64, device=dev())
    w = torch.randn(64, 3, 7, 7, device=dev()) * 0.05
    cols = NV.stem_im2col(x)
    wk = torch.zeros(64, NV.STEM_LDK, device=dev())
    wk[:, :147] = w.permute(0, 2, 3, 1).reshape(64, 147)
    y = NV.gemm_tn(cols, wk.bfloat16())
    ref = F.conv2d(x.bfloat16().float(), w.bfloat16().float(), stride=2, padding=3)
    assert rel_err(y.view(4, 32, 32, 64), nhwc(ref)) < 1e-2


@pytest.mark.parametrize("rows,C", [(8 * 56 * 56, 64), (4 * 28 * 28, 512), (1000, 2048), (37, 24)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_batchnorm(rows, C, relu, res):
    torch.manual_seed(4)
    x = (torch.randn(rows, C, device=dev()) * 2 + 0.5).bfloat16()
    r = torch.randn(rows, C, device=dev()).bfloat16() if res else None
    g = torch.rand(C, device=dev()) + 0.5
    b = torch.randn(C, device=dev())
    rm = torch.zeros(C, device=dev()); rv = torch.ones(C, device=dev())
    xr = x.float().requires_grad_(True)
    rr = r.float().requires_grad_(True) if res else None
    rm_ref = rm.clone(); rv_ref = rv.clone()
    yref = F.batch_norm(xr, rm_ref, rv_ref, g, b, True, 0.1, 1e-5)
    if res:
        yref = yref + rr
    if relu:
        yref = torch.relu(yref)
    sums = NV.bn_stats(x)
    y, mean, invstd = NV.bn_apply(x, r, sums, g, b, rm, rv, 1e-5, 0.1, relu)
    assert rel_err(y, yref) < 1e-2
    assert torch.allclose(rm, rm_ref, atol=1e-3, rtol=1e-3)
    assert torch.allclose(rv, rv_ref, atol=1e-3, rtol=1e-3)
    dy = torch.randn(rows, C, device=dev()).bfloat16()
    gref = g.clone().requires_grad_(True)
    bref = b.clone().requires_grad_(True)
    y2 = F.batch_norm(xr, None, None, gref, bref, True, 0.1, 1e-5)
    if res:
        y2 = y2 + rr
    if relu:
        y2 = torch.relu(y2)
    y2.backward(dy.float())
    dx, dgamma, dbeta, dres = NV.bn_backward(dy, y if res else None, x, mean, invstd, g, relu, res, beta=b)
    # the relu mask is taken from the bf16 output, so allow a little slack
    assert rel_err(dx, xr.grad) < 3e-2
    assert rel_err(dgamma, gref.grad) < 2e-2
    assert rel_err(dbeta, bref.grad) < 2e-2
    if res:
        assert rel_err(dres, rr.grad) < 2e-2
    if relu and res and (rows * C) % 8 == 0:
        # bit-mask variant: the forward writes one bit per element, backward never reads y;
        # the step counter is bumped by the same kernel
        mask = torch.empty(rows * C // 8, device=dev(), dtype=torch.uint8)
        tracked = torch.tensor(41, device=dev(), dtype=torch.int64)
        y3, _, _ = NV.bn_apply(x, r, sums, g, b, rm.clone(), rv.clone(), 1e-5, 0.1, relu, tracked, mask)
        assert torch.equal(y3, y) and int(tracked) == 42
        dx3, dg3, db3, dres3 = NV.bn_backward(dy, None, x, mean, invstd, g, relu, True, beta=b, relu_mask=mask)
        # the mask is taken before the bf16 rounding of y: elements that round to +0 differ, nothing else
        assert rel_err(dx3, dx) < 2e-3 and rel_err(dres3, dres) < 2e-3
        assert rel_err(dg3, dgamma) < 2e-3 and rel_err(db3, dbeta) < 2e-3


@pytest.mark.parametrize("rows,C", [(2048, 4096), (1000, 24), (37, 8), (513, 1000)])
def test_relu_bwd_bias(rows, C):
    torch.manual_seed(21)
    dy = torch.randn(rows, C, device=dev()).bfloat16()
    y = torch.relu(torch.randn(rows, C, device=dev())).bfloat16()
    g, db = NV.relu_bwd_bias(dy, y, True)
    want = dy.float() * (y > 0)
    assert torch.equal(g.float(), want)
    assert torch.allclose(db, want.sum(0), rtol=1e-4, atol=1e-3 * rows ** 0.5)
    g2, db2 = NV.relu_bwd_bias(dy, None, True)
    assert g2 is dy and torch.allclose(db2, dy.float().sum(0), rtol=1e-4, atol=1e-3 * rows ** 0.5)
    g3, db3 = NV.relu_bwd_bias(dy, y, False)
    assert db3 is None and torch.equal(g3, g)


@pytest.mark.parametrize("relu,out_fp32,bias", [(True, False, True), (False, True, True), (True, True, True), (False, False, False)])
def test_linear_autograd_matches_torch(relu, out_fp32, bias):
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    torch.manual_seed(22)
    x = torch.randn(300, 256, device=dev()).bfloat16().float().requires_grad_(True)
    w = (torch.randn(136, 256, device=dev()) / 16).bfloat16().float().requires_grad_(True)
    b = torch.randn(136, device=dev()).requires_grad_(True) if bias else None
    y = Fn.linear(x, w, b, relu, out_fp32)
    assert y.dtype == (torch.float32 if out_fp32 else torch.bfloat16)
    dy = torch.randn_like(y)
    y.backward(dy)
    got = (x.grad.clone(), w.grad.clone(), b.grad.clone() if bias else None)
    x.grad = None; w.grad = None
    if bias:
        b.grad = None
    ref = F.linear(x, w, b)
    if relu:
        ref = torch.relu(ref)
    assert rel_err(y, ref) < 1e-2
    ref.backward(dy.float() * (y.float() > 0) if relu else dy.float())   # same mask as the bf16 output
    assert rel_err(got[0], x.grad) < 1.5e-2
    assert rel_err(got[1], w.grad) < 1.5e-2
    if bias:
        assert rel_err(got[2], b.grad) < 1e-2


def test_pools():
    torch.manual_seed(5)
    x = torch.randn(4, 32, 32, 64, device=dev()).bfloat16()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yref = F.max_pool2d(xr, 3, 2, 1)
    y, arg = NV.maxpool_fwd(x)
    assert torch.equal(y.float(), nhwc(yref))
    dy = torch.randn_like(y)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    dx = NV.maxpool_bwd(dy, arg, x.shape)
    assert rel_err(dx, nhwc(xr.grad)) < 1e-2
    x2 = torch.randn(6, 7, 7, 2048, device=dev()).bfloat16()
    ya = NV.avgpool_fwd(x2)
    assert rel_err(ya, x2.float().mean(dim=(1, 2))) < 1e-2
    dya = torch.randn_like(ya)
    dxa = NV.avgpool_bwd(dya, x2.shape)
    assert rel_err(dxa, (dya.float() / 49)[:, None, None, :].expand(6, 7, 7, 2048)) < 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_losses(dtype):
    torch.manual_seed(6)
    p = torch.randn(120, 1000, device=dev()).to(dtype)
    t = torch.randn(120, 1000, device=dev()).to(dtype)
    pr = p.float().clone().requires_grad_(True)
    lref = F.mse_loss(pr, t.float()); lref.backward()
    loss, grad = NV.mse_loss_fwd_bwd(p, t)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert abs(loss.item() - lref.item()) < tol * max(1, abs(lref.item()))
    assert rel_err(grad, pr.grad) < max(tol, 1e-5)
    # second call on the same scratch must give the same answer (ticket counter reset)
    loss2, _ = NV.mse_loss_fwd_bwd(p, t)
    assert loss2.item() == loss.item()
    s, g = NV.sum_loss_fwd_bwd(p)
    assert abs(s.item() - p.float().sum().item()) < 1e-2 * p.numel() ** 0.5
    assert torch.all(g.float() == 1)
    probs = torch.softmax(torch.randn(120, 1000, device=dev()), -1)
    pr2 = p.float().clone().requires_grad_(True)
    l2 = F.cross_entropy(pr2, probs); l2.backward()
    l, g = NV.ce_prob_fwd_bwd(p, probs)
    assert abs(l.item() - l2.item()) < 1e-3 * abs(l2.item()) + 1e-4
    assert rel_err(g, pr2.grad) < max(tol, 1e-4)
    labels = torch.randint(0, 1000, (120,), device=dev())
    pr3 = p.float().clone().requires_grad_(True)
    l3 = F.cross_entropy(pr3, labels); l3.backward()
    l, g = NV.ce_index_fwd_bwd(p, labels)
    assert abs(l.item() - l3.item()) < 1e-3 * abs(l3.item()) + 1e-4
    assert rel_err(g, pr3.grad) < max(tol, 1e-4)


def test_ce_degenerate_single_class():
    # /root/reference/ddp_gpus.py:37: cross_entropy on [B,1] logits with float targets is identically 0
    out = torch.randn(32, 1, device=dev())
    ys = torch.rand(32, 1, device=dev())
    l, g = NV.ce_prob_fwd_bwd(out, ys)
    assert l.item() == 0.0 and torch.all(g == 0)


def test_layout_helpers():
    x = torch.randn(3, 5, 17, 33, device=dev())
    y = NV.nchw_to_nhwc_bf16(x)
    assert torch.equal(y, x.permute(0, 2, 3, 1).contiguous().bfloat16())
    a = torch.randn(100, 70, device=dev()).bfloat16()
    assert torch.equal(NV.transpose_bf16(a), a.t().contiguous())
    w = torch.randn(1001, device=dev())
    assert torch.equal(NV.cast_bf16(w), w.bfloat16())


def test_multi_copy_kernel_segments():
    """DataParallel's scatter / gather kernel: up to eight (src, dst, bytes) segments in one launch, 16-byte
    vector path and byte path (odd sizes / odd addresses)."""
    import ctypes
    from pytorch_distributed_training_tutorials_b200 import _lib
    dev = torch.device("cuda:0")
    for sizes, skew in (([4096, 16, 1 << 20, 48], 0), ([3, 1001, 17], 0), ([4096, 4096], 4), ([1 << 22], 0)):
        srcs = [torch.randint(0, 255, (n + skew,), dtype=torch.uint8, device=dev)[skew:] for n in sizes]
        dsts = [torch.zeros(n + 32, dtype=torch.uint8, device=dev) for n in sizes]
        nb = (ctypes.c_longlong * len(sizes))(*sizes)
        _lib.check(_lib.lib().b200dt_p2p_multi_copy(_lib.u64_array([t.data_ptr() for t in srcs]),
                                                    _lib.u64_array([t.data_ptr() for t in dsts]), nb, len(sizes),
                                                    _lib.stream_ptr(dev)), "multi_copy")
        torch.cuda.synchronize()
        for s_, d_, n in zip(srcs, dsts, sizes):
            assert torch.equal(d_[:n], s_) and int(d_[n:].sum()) == 0


def test_fused_adam_capturable_matches_torch_and_replays_from_a_graph():
    """Adam with the step count on the device (the reference trains its DataParallel model with Adam,
    01.data_parallel.ipynb:287): eager steps match torch.optim.Adam, and the captured step keeps matching when
    it is replayed -- the bias corrections follow the device-side count, not a value frozen at capture."""
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedAdam
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    shapes = [(257, 33), (1000,), (64, 64), (5,)]
    ps = [torch.randn(s, device=dev, requires_grad=True) for s in shapes]
    qs = [p.detach().clone().requires_grad_(True) for p in ps]
    mine = FusedAdam(ps, lr=1e-2, capturable=True, bf16_shadow=False)
    ref = torch.optim.Adam(qs, lr=1e-2)
    grads = [[torch.randn(s, device=dev) for s in shapes] for _ in range(7)]

    def set_grads(k):
        for p, q, g in zip(ps, qs, grads[k]):
            if p.grad is None:
                p.grad = g.clone(); q.grad = g.clone()
            else:
                p.grad.copy_(g); q.grad.copy_(g)
    for k in range(3):
        set_grads(k)
        mine.step(); ref.step()
    for p, q in zip(ps, qs):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)
    assert float(mine.state[ps[0]]["step"]) == 3.0
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            mine.step()
    torch.cuda.current_stream().wait_stream(s)
    for k in range(3, 7):
        set_grads(k)
        g.replay(); ref.step()
    torch.cuda.synchronize()
    assert float(mine.state[ps[0]]["step"]) == 7.0
    for p, q in zip(ps, qs):
        assert torch.allclose(p, q, rtol=2e-5, atol=2e-6), float((p - q).abs().max())


def test_linear_with_an_output_width_that_is_not_a_multiple_of_8():
    """A 1001-way head: forward on tcgen05, backward on the SIMT kernels -- no cuBLAS fallback."""
    from pytorch_distributed_training_tutorials_b200.ops import nn as bnn
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    lin = bnn.Linear(2048, 1001).to(dev)
    assert not lin.skinny
    ref = torch.nn.Linear(2048, 1001).to(dev)
    ref.load_state_dict(lin.state_dict())
    x = torch.randn(96, 2048, device=dev)
    xr = x.clone().requires_grad_(True); x.requires_grad_(True)
    t = torch.randn(96, 1001, device=dev)
    y = lin(x); yr = ref(xr)
    assert torch.allclose(y, yr, rtol=2e-2, atol=2e-2)
    (y * t).sum().backward(); (yr * t).sum().backward()
    for a, b in ((x.grad, xr.grad), (lin.weight.grad, ref.weight.grad), (lin.bias.grad, ref.bias.grad)):
        assert ((a - b).norm() / b.norm()).item() < 1e-2


@pytest.mark.parametrize("nb,hw", [(4, 224), (3, 128), (2, 32)])
def test_implicit_gemm_stem_matches_fp32_conv(nb, hw):
    """7x7 stride-2 stem without a patch matrix: forward, fused BatchNorm statistics and the weight gradient against
    torch's fp32 convolution on bf16-rounded operands; the weight gradient is bit-identical run to run."""
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    dev = torch.device("cuda:0")
    torch.manual_seed(hw + nb)
    x = torch.randn(nb, 3, hw, hw, device=dev)
    w = torch.randn(64, 7, 7, 3, device=dev) * 0.1
    assert NV.stem_supported(x.shape)
    sums = NV.bn_sums_buffer(64, dev)
    y = NV.stem_conv_fprop(x, w, bn_sums=sums)
    xb, wb = x.bfloat16().float(), w.bfloat16().float()
    ref = torch.nn.functional.conv2d(xb, wb.permute(0, 3, 1, 2), stride=2, padding=3).permute(0, 2, 3, 1)
    assert y.shape == ref.shape
    assert ((y.float() - ref).norm() / ref.norm()).item() < 5e-3
    yf = y.float()
    ref_s = torch.cat([yf.sum((0, 1, 2)), (yf * yf).sum((0, 1, 2))])
    assert torch.allclose(sums[:128], ref_s, rtol=1e-4, atol=1e-2)
    dy = torch.randn_like(y)
    dw = NV.stem_conv_wgrad(x, dy)
    dw2 = NV.stem_conv_wgrad(x, dy)
    assert torch.equal(dw, dw2)
    ref_dw = torch.nn.grad.conv2d_weight(xb, (64, 3, 7, 7), dy.float().permute(0, 3, 1, 2), stride=2,
                                         padding=3).permute(0, 2, 3, 1)
    assert ((dw - ref_dw).norm() / ref_dw.norm()).item() < 5e-3


@pytest.mark.parametrize("C", [64, 256, 1024, 2048])
def test_sums_buffers_zero_chunks_and_the_memset_contract(C):
    """Statistics buffers are slices of zero-filled chunks (one fill instead of a memset per launch) and the
    library is told not to clear them; with the switch off (the C ABI's default) the entry points clear any
    buffer themselves.  Wide layers use fewer partial-sum copies (and 2048 channels none to fold)."""
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    dev = torch.device("cuda:0")
    L = _lib.lib()
    used = NV.bn_replicas_used(C)
    assert used == (NV.bn_replicas() if _lib.is_deterministic() else max(1, min(16, 2048 // C)))
    torch.manual_seed(C)
    x = torch.randn(777, C, device=dev).to(torch.bfloat16)
    xf = x.float()
    ref = torch.cat([xf.sum(0), (xf * xf).sum(0)])
    a = NV.bn_stats(x)
    b = NV.bn_stats(x)
    assert L.b200dt_get_sums_prezeroed() == 1
    assert a.data_ptr() != b.data_ptr() and a.data_ptr() % 256 == 0 and b.data_ptr() % 256 == 0
    assert torch.allclose(a[:2 * C], ref, rtol=1e-4, atol=1e-2) and torch.allclose(b[:2 * C], ref, rtol=1e-4, atol=1e-2)
    # inside a stream capture the chunk (and its fill) belong to the graph: replays start from zeros again
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(g, stream=s):
        c = NV.bn_stats(x)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.allclose(c[:2 * C], ref, rtol=1e-4, atol=1e-2)
    # the C ABI default: a dirty caller-owned buffer is cleared by the call
    L.b200dt_set_sums_prezeroed(0)
    try:
        dirty = torch.full((2 * C * NV.bn_replicas() + 1,), float("nan"), device=dev)
        rc = L.b200dt_bn_stats(x.data_ptr(), dirty.data_ptr(), 777, C, torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        assert torch.allclose(dirty[:2 * C], ref, rtol=1e-4, atol=1e-2)
    finally:
        L.b200dt_set_sums_prezeroed(1)


@pytest.mark.parametrize("nb,H", [(3, 56), (2, 28), (5, 12), (1, 16), (2, 7)])
def test_halo_reuse_3x3_64_to_64(nb, H):
    """3x3 / stride 1 / 64 -> 64 runs in halo mode (one staged patch, nine shifted operand views, resident filter).
    Tiles of 8 x 16 pixels hang over the image edge when H or W is not a multiple: forward, fused statistics, data
    gradient and accumulating data gradient against fp32 torch."""
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    dev = torch.device("cuda:0")
    torch.manual_seed(H * 10 + nb)
    x = torch.randn(nb, H, H, 64, device=dev).to(torch.bfloat16)
    w = (torch.randn(64, 3, 3, 64, device=dev) * 0.06).to(torch.bfloat16)
    dy = torch.randn(nb, H, H, 64, device=dev).to(torch.bfloat16)
    sums = NV.bn_sums_buffer(64, dev)
    y = NV.conv2d_fprop(x, w, bn_sums=sums)
    xf, wf = x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2)
    ref = torch.nn.functional.conv2d(xf, wf, padding=1).permute(0, 2, 3, 1)
    assert ((y.float() - ref).norm() / ref.norm()).item() < 5e-3
    yf = y.float()
    ref_s = torch.cat([yf.sum((0, 1, 2)), (yf * yf).sum((0, 1, 2))])
    assert torch.allclose(sums[:128], ref_s, rtol=1e-4, atol=1e-2)
    dx = NV.conv2d_dgrad(dy, w, x.shape, 3)
    ref_dx = torch.nn.grad.conv2d_input(xf.shape, wf, dy.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    assert ((dx.float() - ref_dx).norm() / ref_dx.norm()).item() < 5e-3
    skip = torch.randn_like(x)
    acc = NV.conv2d_dgrad(dy, w, x.shape, 3, accumulate_into=skip.clone())
    assert ((acc.float() - (skip.float() + ref_dx)).norm() / ref_dx.norm()).item() < 1e-2
