"""Deterministic mode (``set_deterministic`` / B200DT_DETERMINISTIC=1): every sum the kernels normally
combine with atomics is combined in a fixed order, so the same step on the same inputs is bit-identical.
That turns "is this a race or just fp32 ordering noise?" into a yes/no question: any difference in this
mode is a bug.  (profiles/reproducibility.md has the measurements behind the tolerances used elsewhere.)"""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture()
def det():
    from pytorch_distributed_training_tutorials_b200 import _lib
    was = _lib.is_deterministic()
    _lib.set_deterministic(True)
    yield
    _lib.set_deterministic(was)


def _conv_case(NV, H, Cin, Cout, R, s, nb=8):
    dev = torch.device("cuda:0")
    torch.manual_seed(H + Cin + Cout + R + s)
    x = torch.randn(nb, H, H, Cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(Cout, R, R, Cin, device=dev) * (2.0 / (Cin * R * R)) ** 0.5).to(torch.bfloat16)
    dy = torch.randn(nb, H // s, H // s, Cout, device=dev).to(torch.bfloat16)
    return x, w, dy


@pytest.mark.parametrize("shape", [(56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (28, 128, 128, 3, 2), (14, 1024, 256, 1, 1),
                                   (7, 512, 2048, 1, 1)])
def test_kernels_are_bit_identical_run_to_run(det, shape):
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    H, Cin, Cout, R, s = shape
    x, w, dy = _conv_case(NV, *shape)
    dev = x.device
    runs = []
    for rep in range(3):
        sb = NV.bn_sums_buffer(Cout, dev)
        y = NV.conv2d_fprop(x, w, stride=s, bn_sums=sb)
        dw = NV.conv2d_wgrad(x, dy, R, stride=s, out=torch.full((Cout, R, R, Cin), float("nan"), device=dev))
        g = torch.rand(Cout, device=dev) * 0 + 1.25
        b = torch.full((Cout,), 0.1, device=dev)
        x2 = y.reshape(-1, Cout)
        yb, mean, invstd = NV.bn_apply(x2, None, sb, g, b, torch.zeros(Cout, device=dev), torch.ones(Cout, device=dev),
                                       1e-5, 0.1, True)
        dxb, dg, db, _ = NV.bn_backward(dy.reshape(-1, Cout), None, x2, mean, invstd, g, True, False, beta=b)
        st = NV.bn_stats(x2)
        runs.append((y, sb[:2 * Cout].clone(), dw, yb, dxb, dg, db, st[:2 * Cout].clone()))
        torch.empty(1 << 22, device=dev).normal_()     # move the allocator / L2 state between repetitions
    torch.cuda.synchronize()
    names = ["y", "fused sums", "dw", "bn y", "bn dx", "dgamma", "dbeta", "bn_stats sums"]
    for r in runs[1:]:
        for n, a, b_ in zip(names, runs[0], r):
            assert torch.equal(a, b_), (shape, n, float((a.float() - b_.float()).abs().max()))
    # and the fixed-order sums are the right sums
    yf = runs[0][0].float()
    ref = torch.cat([yf.sum((0, 1, 2)), (yf * yf).sum((0, 1, 2))])
    assert torch.allclose(runs[0][1], ref, rtol=1e-4, atol=1e-2)
    assert torch.allclose(runs[0][7], ref, rtol=1e-4, atol=1e-2)


def test_deterministic_results_agree_with_the_default_mode():
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    x, w, dy = _conv_case(NV, 28, 128, 128, 3, 1)
    out = {}
    was = _lib.is_deterministic()
    try:
        for mode in (False, True):
            _lib.set_deterministic(mode)
            sb = NV.bn_sums_buffer(128, x.device)
            y = NV.conv2d_fprop(x, w, bn_sums=sb)
            dw = NV.conv2d_wgrad(x, dy, 3)
            out[mode] = (y, sb[:256].clone(), dw)
    finally:
        _lib.set_deterministic(was)
    assert torch.equal(out[False][0], out[True][0])
    assert torch.allclose(out[False][1], out[True][1], rtol=1e-5, atol=1e-3)
    assert torch.allclose(out[False][2], out[True][2], rtol=1e-4, atol=1e-4)
    # both agree with an fp32 reference of the weight gradient
    ref = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (128, 128, 3, 3), dy.float().permute(0, 3, 1, 2),
                                      padding=1).permute(0, 2, 3, 1)
    assert ((out[True][2] - ref).norm() / ref.norm()).item() < 2e-3


def _step(m, x, y, loss_fn):
    m.zero_grad(set_to_none=True)
    out = m(x)
    loss = loss_fn(out, y)
    loss.backward()
    return out.detach().clone(), loss.detach().clone()


def test_resnet50_step_is_bit_identical(det):
    """Forward output, loss, every gradient and every BatchNorm buffer of two runs of the same ResNet-50 step
    are equal bit for bit -- on the same module and on a deep copy (different addresses for everything)."""
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = resnet50(num_classes=100).to(dev).train()
    m2 = copy.deepcopy(m)
    state0 = copy.deepcopy(m.state_dict())
    x = torch.randn(16, 3, 128, 128, device=dev)
    y = torch.randint(0, 100, (16,), device=dev)
    o1, l1 = _step(m, x, y, Fn.cross_entropy)
    g1 = [p.grad.clone() for p in m.parameters()]
    b1 = [b.clone() for b in m.buffers()]
    m.load_state_dict(state0)
    o2, l2 = _step(m, x, y, Fn.cross_entropy)
    o3, l3 = _step(m2, x, y, Fn.cross_entropy)
    assert torch.equal(o1, o2) and torch.equal(o1, o3)
    assert torch.equal(l1, l2) and torch.equal(l1, l3)
    for (n, p), p2, g in zip(m.named_parameters(), m2.parameters(), g1):
        assert torch.equal(p.grad, g), n
        assert torch.equal(p2.grad, g), n
    for b, b2, ref in zip(m.buffers(), m2.buffers(), b1):
        assert torch.equal(b, ref) and torch.equal(b2, ref)


def test_default_mode_noise_is_ulp_sized_at_the_kernel_level():
    """Without the switch only atomically summed fp32 outputs may move, and only in the last bits; every bf16
    tensor a kernel writes is identical run to run (measured: profiles/reproducibility.md)."""
    from pytorch_distributed_training_tutorials_b200.ops import native as NV
    x, w, dy = _conv_case(NV, 56, 64, 64, 3, 1)
    skip = torch.randn_like(x)
    runs = []
    for _ in range(3):
        sb = NV.bn_sums_buffer(64, x.device)
        runs.append((NV.conv2d_fprop(x, w, bn_sums=sb), NV.conv2d_dgrad(dy, w, x.shape, 3),
                     NV.conv2d_dgrad(dy, w, x.shape, 3, accumulate_into=skip.clone()),
                     sb[:128].clone(), NV.conv2d_wgrad(x, dy, 3)))
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1]) and torch.equal(r[2], runs[0][2])
        assert torch.allclose(r[3], runs[0][3], rtol=2e-6, atol=1e-3)
        assert ((r[4] - runs[0][4]).norm() / runs[0][4].norm()).item() < 2e-6
