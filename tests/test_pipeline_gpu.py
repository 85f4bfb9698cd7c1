"""CUDA-graph replayed pipeline (parallel/graphed_pipeline.py) against the eager micro-batched schedule.
Runs with both stages on one GPU when only one is visible (the schedule and the graphs are the same)."""
import copy

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


def _devs():
    return ["cuda:0", "cuda:1"] if torch.cuda.device_count() >= 2 else ["cuda:0", "cuda:0"]


def _eager_step(stages, devs, x, y, loss_fn, chunks):
    total = 0.0
    for xi, yi in zip(x.chunk(chunks), y.chunk(chunks)):
        a = xi
        for st, d in zip(stages, devs):
            a = st(a.to(d))
        li = loss_fn(a, yi.to(devs[-1])) / chunks
        li.backward()
        total = total + li.detach()
    return total


def test_graphed_pipeline_mlp_matches_eager_over_steps():
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.ops import nn as bnn, functional as Fn
    devs = _devs()
    torch.manual_seed(0)
    s0 = nn.Sequential(bnn.Linear(64, 256, relu=True), bnn.Linear(256, 128, relu=True)).to(devs[0])
    s1 = nn.Sequential(bnn.Linear(128, 64, relu=True), bnn.Linear(64, 32)).to(devs[1])
    r0, r1 = copy.deepcopy(s0), copy.deepcopy(s1)
    x = torch.randn(256, 64, device=devs[0]); y = torch.randn(256, 32, device=devs[1])
    pipe = b2.GraphedPipeline([s0, s1], devs, 4, Fn.mse_loss, x, y)
    assert pipe.launches_per_step == 16
    o = b2.FusedSGD(list(s0.parameters()) + list(s1.parameters()), lr=0.05)
    ro = b2.FusedSGD(list(r0.parameters()) + list(r1.parameters()), lr=0.05)
    losses, ref = [], []
    for it in range(6):
        torch.manual_seed(100 + it)
        xb = torch.randn(256, 64, device=devs[0]); yb = torch.randn(256, 32, device=devs[1])
        # alternate the two ways of clearing gradients: the flat buffers must be re-attached
        if it % 2:
            o.zero_grad(set_to_none=True)
        else:
            pipe.zero_grad()
        losses.append(float(pipe.train_step(xb, yb)))
        o.step()
        ro.zero_grad(set_to_none=True)
        ref.append(float(_eager_step([r0, r1], devs, xb, yb, Fn.mse_loss, 4)))
        ro.step()
    assert ref[-1] < ref[0]
    for a, b in zip(losses, ref):
        assert abs(a - b) <= 2e-3 * abs(b) + 1e-5, (losses, ref)
    for p, q in zip(list(s0.parameters()) + list(s1.parameters()), list(r0.parameters()) + list(r1.parameters())):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-4)


def _cos(a, b):
    a, b = a.flatten().double(), b.flatten().double()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-30))


def test_graphed_pipeline_resnet50_matches_eager():
    """The graph-replayed schedule launches the same kernels on the same numbers as the eager micro-batched
    schedule.  Under fixed-order reductions (deterministic mode) the two are therefore bit-identical: loss,
    boundary activations, every gradient, every BatchNorm buffer.  (Two micro-batches: the accumulation
    order g1 + g0 vs g0 + g1 does not matter for two fp32 terms.)"""
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import PipelineParallelResNet50
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    devs = _devs()
    was = _lib.is_deterministic()
    _lib.set_deterministic(True)
    try:
        torch.manual_seed(0)
        m = PipelineParallelResNet50(split_size=8, num_classes=16, dev0=devs[0], dev1=devs[1], split_after="layer1")
        r, r2 = copy.deepcopy(m), copy.deepcopy(m)
        x = torch.randn(16, 3, 128, 128, device=devs[0]); y = torch.randint(0, 16, (16,), device=devs[1])
        pipe = m.graphed(Fn.cross_entropy, x, y)
        assert pipe.chunks == 2
        # capture must not have disturbed the BatchNorm running statistics or left gradients behind
        for (k, a), (_, b) in zip(m.state_dict().items(), r.state_dict().items()):
            assert torch.equal(a, b), k
        assert all(float(p.grad.abs().sum()) == 0.0 for p in m.parameters())
        loss = pipe.train_step(x, y)
        ref = _eager_step([r.stage0, r.stage1], devs, x, y, Fn.cross_entropy, 2)
        ref2 = _eager_step([r2.stage0, r2.stage1], devs, x, y, Fn.cross_entropy, 2)
        for d in set(devs):
            torch.cuda.synchronize(d)
        # eager twice: identical (this is what makes the comparison below meaningful)
        assert torch.equal(ref, ref2)
        for (n, p), q in zip(r.named_parameters(), r2.parameters()):
            assert torch.equal(p.grad, q.grad), n
        # graph replay vs eager
        assert abs(float(loss) - float(ref)) <= 1e-6 * abs(float(ref))
        for (n, p), q in zip(m.named_parameters(), r.parameters()):
            assert torch.equal(p.grad, q.grad), (n, float((p.grad - q.grad).abs().max()), float(q.grad.abs().max()))
        for (n, b), c in zip(m.named_buffers(), r.named_buffers()):
            assert torch.equal(b, c[1]), n
        assert int(m.bn1.num_batches_tracked) == 2
        with pytest.raises(ValueError):
            pipe.train_step(x[:8], y[:8])
    finally:
        _lib.set_deterministic(was)


def test_block_level_split_and_time_balanced_cut():
    """A cut between two bottleneck blocks computes what the unsplit network computes (bit-identical in
    deterministic mode), and ``balanced()`` picks the cut that minimises the slower stage's measured time."""
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    from pytorch_distributed_training_tutorials_b200.parallel.place import balanced_cuts
    devs = _devs()
    was = _lib.is_deterministic()
    _lib.set_deterministic(True)
    try:
        torch.manual_seed(0)
        whole = resnet50(num_classes=16).to(devs[0])
        sp = ModelParallelResNet50(num_classes=16, dev0=devs[0], dev1=devs[1], split_after="layer2.1")
        sp.load_state_dict(whole.state_dict())
        assert sp.layer2[1].conv1.weight.device == torch.device(devs[0])
        assert sp.layer2[2].conv1.weight.device == torch.device(devs[1])
        x = torch.randn(8, 3, 96, 96, device=devs[0]); y = torch.randint(0, 16, (8,), device=devs[0])
        la = Fn.cross_entropy(whole(x), y); la.backward()
        lb = Fn.cross_entropy(sp(x), y.to(devs[1])); lb.backward()
        assert float(la) == float(lb)
        for (n, p), q in zip(whole.named_parameters(), sp.parameters()):
            assert torch.equal(p.grad.cpu(), q.grad.cpu()), n
    finally:
        _lib.set_deterministic(was)
    bal = ModelParallelResNet50.balanced(torch.randn(16, 3, 96, 96), chunks=2, reps=2, num_classes=16,
                                         dev0=devs[0], dev1=devs[1])
    names = bal.unit_names()
    times = [bal.unit_times_ms[n] for n in names]
    assert all(t > 0 for t in times)
    cut = min(balanced_cuts(times, 2)[0], len(names) - 2)
    assert bal.split_after == names[cut]
    out = bal(torch.randn(4, 3, 96, 96, device=devs[0]))
    assert out.shape == (4, 16) and out.device == torch.device(devs[1])


def test_device_prefetcher_hands_out_the_right_batches():
    from pytorch_distributed_training_tutorials_b200.data import DevicePrefetcher
    batches = [(torch.full((1 << 20,), float(i)).pin_memory(), torch.tensor([i]).pin_memory()) for i in range(7)]
    sums = []
    for x, y in DevicePrefetcher(iter(batches), "cuda:0"):
        assert x.is_cuda and y.is_cuda
        # a long-running consumer: the next upload lands in the other slot while this runs
        acc = x.double()
        for _ in range(20):
            acc = acc * 1.0 + 0.0
        sums.append((acc.sum() / x.numel(), y.clone()))
    assert [round(float(s)) for s, _ in sums] == list(range(7))
    assert [int(y) for _, y in sums] == list(range(7))


def test_trainer_run_batches_graph_matches_step():
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.models.toy import MLP
    torch.manual_seed(0)
    m = MLP(in_features=64, hidden=128, depth=3, out_features=16)
    r = copy.deepcopy(m)
    in_f, out_f = 64, 16
    batches = [(torch.randn(32, in_f).pin_memory(), torch.randint(0, out_f, (32,)).pin_memory()) for _ in range(6)]
    t1 = b2.Trainer(m, None, b2.FusedSGD(m.parameters(), lr=0.05), gpu_id=0, wrap_ddp=False, cuda_graph=True)
    t2 = b2.Trainer(r, None, b2.FusedSGD(r.parameters(), lr=0.05), gpu_id=0, wrap_ddp=False, cuda_graph=False)
    got = t1.run_batches(iter(batches))
    ref = [t2.step(x, y) for x, y in batches]
    assert len(got) == 6
    for a, b in zip(got, ref):
        assert abs(a - b) <= 2e-3 * abs(b) + 1e-5, (got, ref)
