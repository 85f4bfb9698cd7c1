"""Single-process multi-GPU tests (need >= 2 GPUs): DataParallel vs nn.DataParallel, naive and
pipelined model parallelism, multi-device optimizer, auto placement on real devices."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")


@needs2
def test_data_parallel_matches_torch(capsys):
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import DataParallel
    from pytorch_distributed_training_tutorials_b200.models.toy import SampleModel
    n = min(_ngpu(), 4)
    torch.manual_seed(0)
    base = SampleModel(32, 2, verbose=True).to("cuda:0")
    ref_m = copy.deepcopy(base); ref_m.verbose = False
    mine = DataParallel(base, device_ids=list(range(n)))
    ref = torch.nn.DataParallel(ref_m, device_ids=list(range(n)))
    o1 = torch.optim.Adam(mine.parameters(), lr=1e-3); o2 = torch.optim.Adam(ref.parameters(), lr=1e-3)
    for it in range(5):
        x = torch.randn(32, 32, device="cuda:0")
        o1.zero_grad(); out = mine(x); out.sum().backward(); o1.step()
        o2.zero_grad(); out2 = ref(x); out2.sum().backward(); o2.step()
        assert out.shape == (32, 2) and out.device == torch.device("cuda:0")
        assert torch.allclose(out, out2, atol=1e-5)
    printed = capsys.readouterr().out
    per = 32 // n
    assert printed.count(f"In Model: input size torch.Size([{per}, 32])") == 5 * n
    for p, q in zip(base.parameters(), ref_m.parameters()):
        assert torch.allclose(p, q, atol=1e-5)
    # gather fused with the sum loss gives the same loss and the same gradients
    base.verbose = False
    o1.zero_grad(); s, out = mine.forward_sum(x); s.backward()
    g1 = [p.grad.clone() for p in base.parameters()]
    o1.zero_grad(); mine(x).sum().backward()
    for a, p in zip(g1, base.parameters()):
        assert torch.allclose(a, p.grad, atol=1e-4)
    assert torch.allclose(s, mine(x).sum(), atol=1e-3)


@needs2
def test_data_parallel_batchnorm_buffers_and_big_model():
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import DataParallel
    n = min(_ngpu(), 4)
    torch.manual_seed(1)
    m = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.BatchNorm1d(256), torch.nn.ReLU(), torch.nn.Linear(256, 7)).to("cuda:0")
    r = copy.deepcopy(m)
    mine = DataParallel(m, device_ids=list(range(n))); ref = torch.nn.DataParallel(r, device_ids=list(range(n)))
    x = torch.randn(64, 64, device="cuda:0"); y = torch.randn(64, 7, device="cuda:0")
    F.mse_loss(mine(x), y).backward(); F.mse_loss(ref(x), y).backward()
    for p, q in zip(m.parameters(), r.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5), (p.grad - q.grad).abs().max()
    # replica 0 aliases the original buffers: running stats were updated in place, like torch
    assert torch.allclose(m[1].running_mean, r[1].running_mean, atol=1e-6)
    assert int(m[1].num_batches_tracked) == 1


@pytest.mark.skipif(_ngpu() < 3, reason="the two-level broadcast / reduce needs >= 3 GPUs")
def test_data_parallel_two_level_transfers_match_torch():
    """>= 3 devices and >= 1 MB of parameters: scatter + all-gather broadcast, reduce-scatter + gather."""
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import DataParallel
    from pytorch_distributed_training_tutorials_b200.parallel import comm
    n = min(_ngpu(), 8)
    torch.manual_seed(4)
    m = torch.nn.Sequential(torch.nn.Linear(96, 1024), torch.nn.ReLU(), torch.nn.Linear(1024, 1000), torch.nn.ReLU(),
                            torch.nn.Linear(1000, 5)).to("cuda:0")
    assert sum(p.numel() for p in m.parameters()) >= comm.TWO_LEVEL_MIN_ELEMS
    r = copy.deepcopy(m)
    mine = DataParallel(m, device_ids=list(range(n))); ref = torch.nn.DataParallel(r, device_ids=list(range(n)))
    o1 = torch.optim.SGD(m.parameters(), lr=0.05); o2 = torch.optim.SGD(r.parameters(), lr=0.05)
    for it in range(4):
        x = torch.randn(16 * n, 96, device="cuda:0"); y = torch.randn(16 * n, 5, device="cuda:0")
        o1.zero_grad(); F.mse_loss(mine(x), y).backward(); o1.step()
        o2.zero_grad(); F.mse_loss(ref(x), y).backward(); o2.step()
    for p, q in zip(m.parameters(), r.parameters()):
        assert torch.allclose(p, q, atol=2e-5, rtol=1e-4), (p - q).abs().max()
    # the replicas really received the parameters (not just replica 0)
    reps = comm.broadcast_coalesced([p.detach() for p in m.parameters()], [torch.device("cuda", i) for i in range(n)])
    for d in range(1, n):
        for a, b in zip(reps[d], m.parameters()):
            assert a.device.index == d and torch.equal(a.cpu(), b.detach().cpu())


@needs2
def test_toy_model_parallel_and_multi_device_optimizer():
    from pytorch_distributed_training_tutorials_b200.models.toy import ToyModel
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    torch.manual_seed(2)
    m = ToyModel("cuda:0", "cuda:1")
    assert m.net1.weight.device == torch.device("cuda:0") and m.net2.weight.device == torch.device("cuda:1")
    ref = copy.deepcopy(m)
    opt = FusedSGD(m.parameters(), lr=1e-3, bf16_shadow=False); ropt = torch.optim.SGD(ref.parameters(), lr=1e-3)
    for _ in range(3):
        x = torch.randn(20, 10000); labels = torch.randn(20, 5).to("cuda:1")
        opt.zero_grad(); out = m(x); assert out.device == torch.device("cuda:1")
        F.mse_loss(out, labels).backward(); opt.step()
        ropt.zero_grad(); F.mse_loss(ref(x), labels).backward(); ropt.step()
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-6)


@needs2
def test_resnet50_split_naive_and_pipelined():
    from pytorch_distributed_training_tutorials_b200 import _lib
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
    from pytorch_distributed_training_tutorials_b200.parallel.model_parallel import model_size
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    torch.manual_seed(3)
    single = resnet50(num_classes=1000).to("cuda:0")
    mp = ModelParallelResNet50(num_classes=1000)
    pp = PipelineParallelResNet50(split_size=8, num_classes=1000)
    assert model_size(mp) == model_size(single) == 25557032
    mp.load_state_dict(single.state_dict()); pp.load_state_dict(single.state_dict())
    assert mp.layer2[0].conv1.weight.device == torch.device("cuda:0")
    assert mp.layer3[0].conv1.weight.device == torch.device("cuda:1") and mp.fc.weight.device == torch.device("cuda:1")
    x = torch.randn(32, 3, 128, 128)
    labels = torch.zeros(32, 1000).scatter_(1, torch.randint(0, 1000, (32, 1)), 1)
    # The split model runs the same kernels on the same numbers, only on two devices.  With fixed-order
    # reductions that makes it BIT-IDENTICAL to the single-GPU model: outputs, gradients, BN buffers.
    was = _lib.is_deterministic()
    _lib.set_deterministic(True)
    try:
        out_s = single(x.to("cuda:0"))
        out_m = mp(x)
        assert out_m.device == torch.device("cuda:1")
        assert torch.equal(out_s.cpu(), out_m.cpu())
        Fn.mse_loss(out_m, labels.to("cuda:1")).backward()
        Fn.mse_loss(out_s, labels.to("cuda:0")).backward()
    finally:
        _lib.set_deterministic(was)
    for (n, p), q in zip(mp.named_parameters(), single.parameters()):
        assert p.grad.device == p.device, n
        assert torch.equal(p.grad.cpu(), q.grad.cpu()), n
    for (n, b), c in zip(mp.named_buffers(), single.buffers()):
        assert torch.equal(b.cpu(), c.cpu()), n
    # pipelined: BN sees micro-batches, so only check shape, devices, and that it trains
    out_p = pp(x)
    assert out_p.shape == (32, 1000) and out_p.device == torch.device("cuda:1")
    Fn.mse_loss(out_p, labels.to("cuda:1")).backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in pp.parameters())


@needs2
def test_auto_place_dispatch_runs_across_devices():
    from pytorch_distributed_training_tutorials_b200.models.llama_like import LlamaLikeForCausalLM
    from pytorch_distributed_training_tutorials_b200.parallel.place import auto_place, dispatch, placement_table
    n = min(_ngpu(), 4)
    m = LlamaLikeForCausalLM(vocab=256, dim=64, layers=8, heads=4, hidden=128)
    ref = copy.deepcopy(m).to("cuda:0")
    pl = auto_place(m, list(range(n)), blocks=m.placement_blocks())
    dispatch(m, pl)
    devs = {str(d) for _, _, _, d in placement_table(m)}
    assert len(devs) == n
    ids = torch.randint(0, 256, (2, 16), device="cuda:0")
    out = m(ids)
    assert out.device.index == n - 1
    assert torch.allclose(out.cpu(), ref(ids).cpu(), atol=1e-4)
    out.float().mean().backward()
    assert all(p.grad is not None for p in m.parameters())


@needs2
def test_graphed_data_parallel_matches_single_device_training():
    """Persistent replicas + one CUDA graph per device: same losses and parameters as training the
    whole mini-batch on one GPU (mean-reduced loss), including the per-step state broadcast."""
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.models.toy import MLP
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    n = min(_ngpu(), 4)
    torch.manual_seed(6)
    m = MLP(in_features=64, hidden=256, depth=3, out_features=16).to("cuda:0")
    r = copy.deepcopy(m)
    o = b2.FusedSGD(m.parameters(), lr=0.05); ro = b2.FusedSGD(r.parameters(), lr=0.05)
    B = 32 * n
    x = torch.randn(B, 64, device="cuda:0"); y = torch.randn(B, 16, device="cuda:0")
    gdp = b2.GraphedDataParallel(m, list(range(n)), Fn.mse_loss, x, y)
    assert gdp.launches_per_step == n
    got, ref = [], []
    for it in range(5):
        torch.manual_seed(50 + it)
        xb = torch.randn(B, 64, device="cuda:0"); yb = torch.randn(B, 16, device="cuda:0")
        o.zero_grad()                                   # set_to_none: the flat views must be re-attached
        got.append(float(gdp.train_step(xb, yb)))
        o.step()
        ro.zero_grad()
        l = Fn.mse_loss(r(xb), yb); l.backward(); ro.step()
        ref.append(float(l))
    for a, b in zip(got, ref):
        assert abs(a - b) <= 2e-3 * abs(b) + 1e-5, (got, ref)
    for p, q in zip(m.parameters(), r.parameters()):
        assert torch.allclose(p, q, rtol=2e-3, atol=2e-4), (p - q).abs().max()
    # replicas carry the root's current parameters after the next broadcast
    gdp.train_step(xb, yb)
    for d in range(1, n):
        for p0, pd in zip(m.parameters(), gdp.replicas[d].parameters()):
            assert pd.device.index == d
    with pytest.raises(ValueError):
        gdp.train_step(xb[:8], yb[:8])


@needs2
def test_data_parallel_replica_cache_and_kernel_scatter_gather():
    """Parameters are re-broadcast only after something wrote them (version counters); scatter and gather are one
    kernel each and agree with torch's DataParallel, for even and ragged batches."""
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.models.toy import MLP
    n = min(_ngpu(), 8)
    torch.manual_seed(0)
    m = MLP(64, 128, 2, 16).to("cuda:0")
    ref = copy.deepcopy(m)
    dp = b2.DataParallel(m, device_ids=list(range(n)))
    rdp = torch.nn.DataParallel(ref, device_ids=list(range(n)))
    opt = b2.FusedSGD(m.parameters(), lr=0.05, bf16_shadow=False)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05)
    for B in (8 * n, 8 * n + 3, n - 1 if n > 2 else 3):
        x = torch.randn(B, 64, device="cuda:0")
        with torch.no_grad():
            a = dp(x); b = dp(x)                     # second forward: nothing was written -> no broadcast
        assert torch.equal(a, b)
        c = rdp(x)
        assert a.shape == c.shape and torch.allclose(a.float(), c.float(), atol=3e-2, rtol=3e-2)
    pc = dp._replica_cache["params"]
    assert pc["hits"] >= 3 and pc["misses"] >= 1
    misses = pc["misses"]
    x = torch.randn(8 * n, 64, device="cuda:0"); y = torch.randn(8 * n, 16, device="cuda:0")
    seen = []
    for it in range(3):
        opt.zero_grad(); torch.nn.functional.mse_loss(dp(x).float(), y).backward(); opt.step()
        ropt.zero_grad(); torch.nn.functional.mse_loss(rdp(x), y).backward(); ropt.step()
        seen.append(pc["misses"])
    # every optimizer step invalidated the replicas: each forward after one is a miss.  (The first forward of the
    # loop is a miss only if the last ragged batch above used fewer devices: the device set is part of the key.)
    assert seen[1] == seen[0] + 1 and seen[2] == seen[1] + 1 and seen[0] in (misses, misses + 1)
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=2e-2, rtol=2e-2)
    # gather of scalars (one per replica) and the gradient scatter behind it
    class S(torch.nn.Module):
        def __init__(self):
            super().__init__(); self.w = torch.nn.Parameter(torch.ones(4))
        def forward(self, x):
            return (x * self.w).sum()
    sm = S().to("cuda:0"); sdp = b2.DataParallel(sm, device_ids=list(range(n)))
    xs = torch.randn(4 * n, 4, device="cuda:0")
    out = sdp(xs)
    assert out.shape == (n,)
    out.sum().backward()
    assert torch.allclose(sm.w.grad, xs.sum(0), atol=1e-4)
