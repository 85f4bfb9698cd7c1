"""The tutorial's degenerate Linear layers on the fp32 SIMT kernel (csrc/skinny.cu) against torch.
Collected after the other kernel tests on purpose: the element code of these kernels is unit-tested
on the host (tests/test_plumbing_cpu.py), this file is their first contact with a device."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,K,N", [(32, 20, 1), (8, 32, 2), (20, 10000, 10), (20, 10, 5), (2048, 20, 1), (7, 3, 9)])
def test_skinny_linear_matches_torch(B, K, N):
    from pytorch_distributed_training_tutorials_b200.ops import nn as bnn
    torch.manual_seed(K)
    m = bnn.Linear(K, N).cuda()
    assert m.skinny
    ref = torch.nn.Linear(K, N).cuda()
    ref.load_state_dict(m.state_dict())
    x = torch.randn(B, K, device="cuda", requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    dy = torch.randn(B, N, device="cuda")
    y = m(x); yr = ref(xr)
    assert y.dtype == torch.float32 and torch.allclose(y, yr, rtol=1e-4, atol=1e-4)
    y.backward(dy); yr.backward(dy)
    assert torch.allclose(x.grad, xr.grad, rtol=1e-4, atol=1e-4)
    assert torch.allclose(m.weight.grad, ref.weight.grad, rtol=1e-4, atol=1e-3)
    assert torch.allclose(m.bias.grad, ref.bias.grad, rtol=1e-4, atol=1e-3)


def test_toy_models_train_without_a_library_gemm():
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.models.toy import ToyLinear, SampleModel
    torch.manual_seed(0)
    m = ToyLinear().cuda()
    opt = b2.FusedSGD(m.parameters(), lr=0.05)
    x = torch.randn(32, 20, device="cuda"); t = x.sum(1, keepdim=True) * 0.1
    before = _lib.launch_count
    losses = []
    for _ in range(30):
        opt.zero_grad(); l = F.mse_loss(m(x), t); l.backward(); opt.step(); losses.append(float(l))
    assert losses[-1] < 0.5 * losses[0]
    assert _lib.launch_count - before >= 60          # forward + backward + optimizer launches are ours
    s = SampleModel(32, 2, verbose=False).cuda()
    assert s(torch.randn(8, 32, device="cuda")).shape == (8, 2)
