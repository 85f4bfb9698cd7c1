"""DDP plumbing on CPU: 2 real processes over gloo (BASELINE.json config 1)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist

import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200 import group
from pytorch_distributed_training_tutorials_b200.launch.spawn import spawn, ProcessRaisedException
from pytorch_distributed_training_tutorials_b200.parallel.ddp import DistributedDataParallel as DDP


def _setup(rank, world, port):
    group.init(backend="gloo", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)


def _grad_equality(rank, world, port, out_dir):
    _setup(rank, world, port)
    torch.manual_seed(100 + rank)                       # different init per rank: ctor must fix that
    model = torch.nn.Sequential(torch.nn.Linear(20, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1))
    ddp = DDP(model)
    ps = [p.detach().clone() for p in model.parameters()]
    gathered = [torch.zeros_like(ps[0]) for _ in range(world)]
    dist.all_gather(gathered, ps[0])
    assert all(torch.equal(g, gathered[0]) for g in gathered), "rank-0 broadcast must make params identical"
    assert list(ddp.state_dict().keys())[0].startswith("module.")
    opt = torch.optim.SGD(ddp.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(7)
    X = torch.randn(3, world * 8, 20, generator=g); Y = torch.randn(3, world * 8, 1, generator=g)
    ref = torch.nn.Sequential(torch.nn.Linear(20, 16), torch.nn.ReLU(), torch.nn.Linear(16, 1))
    ref.load_state_dict(model.state_dict())
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1)
    for it in range(3):
        xs = X[it, rank * 8:(rank + 1) * 8]; ys = Y[it, rank * 8:(rank + 1) * 8]
        opt.zero_grad()
        torch.nn.functional.mse_loss(ddp(xs), ys).backward()
        # single-process oracle on the full batch == mean of per-rank gradients
        ropt.zero_grad()
        torch.nn.functional.mse_loss(ref(X[it]), Y[it]).backward()
        for p, q in zip(model.parameters(), ref.parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-6), (it, (p.grad - q.grad).abs().max())
        opt.step(); ropt.step()
    info = ddp.get_logging_data()
    assert info["rebuilt"] and info["iterations"] == 3
    # no_sync: gradients stay local, then one synchronised backward reduces the accumulated sum
    opt.zero_grad()
    with ddp.no_sync():
        torch.nn.functional.mse_loss(ddp(X[0, rank * 8:(rank + 1) * 8]), Y[0, rank * 8:(rank + 1) * 8]).backward()
    local = [p.grad.clone() for p in model.parameters()]
    gl = [torch.zeros_like(local[0]) for _ in range(world)]
    dist.all_gather(gl, local[0])
    assert not torch.equal(gl[0], gl[1]), "no_sync must not all-reduce"
    torch.nn.functional.mse_loss(ddp(X[1, rank * 8:(rank + 1) * 8]), Y[1, rank * 8:(rank + 1) * 8]).backward()
    gl2 = [torch.zeros_like(local[0]) for _ in range(world)]
    dist.all_gather(gl2, next(model.parameters()).grad.clone())
    assert torch.allclose(gl2[0], gl2[1])
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def _find_port():
    return group.find_free_port()


def test_ddp_gloo_grad_equality(tmp_path):
    spawn(_grad_equality, args=(2, _find_port(), str(tmp_path)), nprocs=2)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _toy_reference_run(rank, world, port, out_dir):
    """The reference workload: Linear(20,1), 2048 samples, batch 32, degenerate cross-entropy."""
    from torch.utils.data import DataLoader
    from pytorch_distributed_training_tutorials_b200.data import DistributedSampler, MyTrainDataset
    from pytorch_distributed_training_tutorials_b200.trainer import Trainer
    _setup(rank, world, port)
    ds = MyTrainDataset(2048)
    loader = DataLoader(ds, batch_size=32, pin_memory=False, shuffle=False, sampler=DistributedSampler(ds))
    model = torch.nn.Linear(20, 1)
    opt = torch.optim.SGD(model.parameters(), lr=1e-2)
    tr = Trainer(model, loader, opt, gpu_id=rank)
    w0 = model.weight.detach().clone()
    import io, contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        tr.train(2)
    lines = buf.getvalue().strip().splitlines()
    assert lines[0] == f"[GPU: {rank} Epoch: 0, Batch size: 32 | Steps 32]", lines
    assert len(lines) == 2 and tr.global_step == 64
    # F.cross_entropy on [B,1] logits with float targets is identically zero: weights never move
    assert torch.equal(model.weight.detach(), w0)
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def test_toy_ddp_steps_and_banner(tmp_path):
    spawn(_toy_reference_run, args=(2, _find_port(), str(tmp_path)), nprocs=2)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _raiser(rank, world, port):
    _setup(rank, world, port)
    if rank == 1:
        raise ValueError("boom on rank 1")
    dist.barrier()          # rank 0 would hang here forever without sibling teardown
    group.destroy()


def test_spawn_propagates_child_error():
    with pytest.raises(ProcessRaisedException) as ei:
        spawn(_raiser, args=(2, _find_port()), nprocs=2)
    assert ei.value.rank == 1 and "boom on rank 1" in str(ei.value)


def _mismatch(rank, world, port):
    _setup(rank, world, port)
    model = torch.nn.Linear(20, 1 + rank)
    DDP(model)


def test_ddp_detects_shape_mismatch():
    with pytest.raises(ProcessRaisedException) as ei:
        spawn(_mismatch, args=(2, _find_port()), nprocs=2)
    assert "same model across all ranks" in str(ei.value)


def _fault_injected_trainer(rank, world, port):
    """Trainer under DDP with B200DT_FAULT_INJECT: rank 1 dies at global step 2; with B200DT_DEBUG the
    bucket handshake runs before every all-reduce; with B200DT_TRACE the ranges are recorded."""
    os.environ["B200DT_FAULT_INJECT"] = "rank=1,step=2"
    os.environ["B200DT_DEBUG"] = "1"
    _setup(rank, world, port)
    from torch.utils.data import DataLoader
    ds = b2.MyTrainDataset(64)
    loader = DataLoader(ds, batch_size=8, shuffle=False, sampler=b2.DistributedSampler(ds))
    model = torch.nn.Linear(20, 1)
    tr = b2.Trainer(model, loader, torch.optim.SGD(model.parameters(), lr=0.01), gpu_id=rank,
                    loss_fn=torch.nn.functional.mse_loss)
    assert tr.model._debug
    tr.train(2)             # 4 steps per epoch: the fault fires in the first epoch
    group.destroy()


def test_fault_injection_is_reported_with_rank_and_siblings_are_torn_down():
    with pytest.raises(ProcessRaisedException) as ei:
        spawn(_fault_injected_trainer, args=(2, _find_port()), nprocs=2)
    assert ei.value.rank == 1 and "fault injected on rank 1 at step 2" in str(ei.value)


def _handshake_mismatch(rank, world, port):
    os.environ["B200DT_DEBUG"] = "1"
    _setup(rank, world, port)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Linear(8, 1))
    ddp = DDP(model)
    # desynchronise the ranks on purpose: rank 1 believes it already issued one more collective
    if rank == 1:
        ddp._debug_seq += 1
    torch.nn.functional.mse_loss(ddp(torch.randn(4, 8)), torch.randn(4, 1)).backward()


def test_debug_handshake_catches_desynchronised_collectives():
    with pytest.raises(ProcessRaisedException) as ei:
        spawn(_handshake_mismatch, args=(2, _find_port()), nprocs=2)
    assert "collective mismatch" in str(ei.value)


def _resnet_ddp(rank, world, port, out_dir):
    """ResNet-50 (the modules' CPU reference path) under DDP over gloo: rank 0's parameters and BatchNorm buffers are
    broadcast, gradients are averaged over the ranks, replicas stay identical step after step."""
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    _setup(rank, world, port)
    torch.manual_seed(100 + rank)                       # different init per rank
    model = resnet50(num_classes=8)
    ddp = DDP(model)
    opt = torch.optim.SGD(ddp.parameters(), lr=0.01)
    g = torch.Generator().manual_seed(50 + rank)        # different data per rank

    def same_everywhere(t):
        got = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(got, t.detach().contiguous())
        return all(torch.equal(x, got[0]) for x in got)

    assert all(same_everywhere(p) for p in list(model.parameters())[::20])
    for it in range(2):
        x = torch.randn(2, 3, 32, 32, generator=g); y = torch.randint(0, 8, (2,), generator=g)
        opt.zero_grad()
        torch.nn.functional.cross_entropy(ddp(x), y).backward()
        assert all(same_everywhere(p.grad) for p in list(model.parameters())[::20]), it
        opt.step()
    assert all(same_everywhere(p) for p in model.parameters())
    # buffers: every forward starts from rank 0's running statistics (broadcast_buffers=True, the reference's default)
    ddp(torch.randn(2, 3, 32, 32, generator=g))
    model.eval()
    with torch.no_grad():
        ddp(torch.randn(2, 3, 32, 32, generator=g))      # eval forward: broadcast, no update
    assert same_everywhere(model.bn1.running_mean) and same_everywhere(model.layer4[2].bn3.running_var)
    assert int(model.bn1.num_batches_tracked) == 3
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def test_resnet50_ddp_over_gloo_keeps_replicas_identical(tmp_path):
    spawn(_resnet_ddp, args=(2, _find_port(), str(tmp_path)), nprocs=2)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
