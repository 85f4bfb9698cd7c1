"""ResNet-50 on the CPU reference path of the modules (plain torch ops behind the same NHWC interface): topology,
parameter mapping and the split / pipelined variants, checked without a GPU.  fp64 removes conditioning from the
comparison (a random-init 50-layer BatchNorm network amplifies fp32 rounding to tens of percent in some gradients;
profiles/reproducibility.md), so every gradient must agree with torchvision's to rounding error."""
import pytest
import torch

torchvision = pytest.importorskip("torchvision")


def _pair(num_classes=10):
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    torch.manual_seed(0)
    ref = torchvision.models.resnet50(num_classes=num_classes).double()
    mine = resnet50(num_classes=num_classes).double()
    assert not mine.load_state_dict(ref.state_dict()).missing_keys
    return ref, mine


def _nchw(g):
    return g.permute(0, 3, 1, 2) if g.dim() == 4 else g


def test_cpu_resnet50_matches_torchvision_forward_backward_and_buffers():
    ref, mine = _pair()
    x = torch.randn(2, 3, 64, 64, dtype=torch.float64)
    y = torch.randint(0, 10, (2,))
    ref.train(); mine.train()
    la = torch.nn.functional.cross_entropy(ref(x), y)
    lb = torch.nn.functional.cross_entropy(mine(x), y)
    assert abs(float(la.detach()) - float(lb.detach())) < 1e-10
    la.backward(); lb.backward()
    rg = dict(ref.named_parameters())
    for n, p in mine.named_parameters():
        g, r = _nchw(p.grad), rg[n].grad
        assert float((g - r).abs().max()) <= 1e-9 * float(r.abs().max()) + 1e-14, n
    for (n, b), (_, c) in zip(mine.named_buffers(), ref.named_buffers()):
        assert torch.allclose(b.double(), c.double(), atol=1e-12), n
    assert int(mine.bn1.num_batches_tracked) == 1
    ref.eval(); mine.eval()
    with torch.no_grad():
        assert torch.allclose(ref(x), mine(x), atol=1e-10)
    # and the exported state_dict is torchvision's again (OIHW filters)
    sd = mine.state_dict()
    assert set(sd) == set(ref.state_dict())
    assert sd["layer1.0.conv2.weight"].shape == ref.state_dict()["layer1.0.conv2.weight"].shape


def test_cpu_split_and_pipelined_resnet_equal_the_whole_model():
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
    _, whole = _pair()
    x = torch.randn(4, 3, 64, 64, dtype=torch.float64)
    whole.eval()
    with torch.no_grad():
        want = whole(x)
    for cls, kw in ((ModelParallelResNet50, {}), (PipelineParallelResNet50, {"split_size": 2})):
        for cut in ("layer2", "layer1.1", "layer3.4"):
            sp = cls(num_classes=10, dev0="cpu", dev1="cpu", split_after=cut, **kw).double()
            sp.load_state_dict(whole.state_dict())
            sp.eval()                                  # eval: micro-batching does not change BatchNorm statistics
            with torch.no_grad():
                got = sp(x)
            assert torch.allclose(got, want, atol=1e-10), (cls.__name__, cut)
    # training mode: the naive split is the same computation as the whole model, gradient for gradient
    whole.train()
    sp = ModelParallelResNet50(num_classes=10, dev0="cpu", dev1="cpu", split_after="layer2.2").double()
    sp.load_state_dict(whole.state_dict())
    y = torch.randint(0, 10, (4,))
    torch.nn.functional.cross_entropy(whole(x), y).backward()
    torch.nn.functional.cross_entropy(sp(x), y).backward()
    for (n, p), q in zip(whole.named_parameters(), sp.parameters()):
        assert float((p.grad - q.grad).abs().max()) <= 1e-9 * float(p.grad.abs().max()) + 1e-14, n


def test_cpu_time_balanced_cut_picks_a_valid_unit():
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50
    from pytorch_distributed_training_tutorials_b200.parallel.place import balanced_cuts
    torch.manual_seed(0)
    m = ModelParallelResNet50.balanced(torch.randn(2, 3, 64, 64), chunks=1, reps=1, num_classes=10,
                                       dev0="cpu", dev1="cpu")
    names = m.unit_names()
    times = [m.unit_times_ms[n] for n in names]
    assert len(times) == 18 and all(t > 0 for t in times)
    assert m.split_after == names[min(balanced_cuts(times, 2)[0], len(names) - 2)]
    assert m(torch.randn(2, 3, 64, 64)).shape == (2, 10)
