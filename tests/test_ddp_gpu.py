"""Multi-GPU tests (need >= 2 GPUs on the box): fused peer-memory collectives vs NCCL, our DDP vs torch DDP."""
import os

import pytest
import torch
import torch.distributed as dist

pytestmark = pytest.mark.gpu

from pytorch_distributed_training_tutorials_b200 import group
from pytorch_distributed_training_tutorials_b200.launch.spawn import spawn


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")


def _world():
    n = _ngpu()
    return 8 if n >= 8 else 4 if n >= 4 else 2


def _collectives(rank, world, port, out_dir):
    from pytorch_distributed_training_tutorials_b200.ops import collectives as coll
    st = group.init(backend="nccl", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    dev = st.device
    arena = coll.SymmetricArena(64 << 20, dev)
    modes = ["oneshot", "twoshot"] + (["nvls"] if arena.has_multicast else [])
    torch.manual_seed(1000 + rank)
    sizes = [1, 7, 21, 256, 1000, 4099, 65536, 300001, 2_000_003]
    for mode in modes:
        for wire_bf16 in (False, True):
            for rep in range(3):
                ts = [torch.randn(n, device=dev) for n in sizes]
                ref = [t.clone() for t in ts]
                for r in ref:
                    dist.all_reduce(r)
                    r.div_(world)
                coll.all_reduce_(ts, arena, mode=mode, scale=1.0 / world, wire_bf16=wire_bf16)
                torch.cuda.synchronize(dev)
                arena.check_status()
                tol = 2e-2 if wire_bf16 else 1e-5
                for a, b in zip(ts, ref):
                    err = (a - b).abs().max().item()
                    assert err <= tol * max(1.0, b.abs().max().item()), (mode, wire_bf16, a.numel(), err)
    # results must be bit-identical on every rank (fixed summation order)
    t = torch.randn(100_003, device=dev)
    coll.all_reduce_([t], arena, mode="twoshot")
    g = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(g, t)
    assert all(torch.equal(g[0], x) for x in g)
    # stress: many back-to-back launches with skew between ranks
    for it in range(200):
        if (it + rank) % 7 == 0:
            torch.cuda._sleep(200_000)
        x = torch.full((4096 + it,), float(rank + 1), device=dev)
        coll.all_reduce_([x], arena, mode=modes[it % len(modes)])
        if it % 50 == 49:
            assert torch.all(x == world * (world + 1) / 2), it
    arena.check_status()
    # broadcast of mixed dtypes through the raw-word path
    a = torch.randn(1001, device=dev); b = torch.arange(10, device=dev, dtype=torch.int32) * (rank + 1)
    coll.broadcast_([a, b.view(torch.int32)], arena, root=0)
    ga = [torch.empty_like(a) for _ in range(world)]
    dist.all_gather(ga, a)
    assert all(torch.equal(ga[0], x) for x in ga)
    assert torch.equal(b.cpu(), torch.arange(10, dtype=torch.int32))
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write(",".join(modes))


@needs2
def test_fused_collectives_match_nccl(tmp_path):
    w = _world()
    spawn(_collectives, args=(w, group.find_free_port(), str(tmp_path)), nprocs=w)
    assert all((tmp_path / f"ok{r}").exists() for r in range(w))


def _ddp_vs_torch(rank, world, port, out_dir, mode, comm_dtype):
    from pytorch_distributed_training_tutorials_b200.parallel.ddp import DistributedDataParallel as DDP
    st = group.init(backend="nccl", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    dev = st.device
    torch.manual_seed(5)

    def make():
        torch.manual_seed(5)
        return torch.nn.Sequential(torch.nn.Linear(64, 512), torch.nn.ReLU(), torch.nn.Linear(512, 2048),
                                   torch.nn.ReLU(), torch.nn.BatchNorm1d(2048), torch.nn.Linear(2048, 10)).to(dev)
    mine = DDP(make(), device_ids=[dev.index], allreduce=mode, comm_dtype=comm_dtype, bucket_cap_mb=1)
    ref = torch.nn.parallel.DistributedDataParallel(make(), device_ids=[dev.index], bucket_cap_mb=1)
    o1 = torch.optim.SGD(mine.parameters(), lr=0.05); o2 = torch.optim.SGD(ref.parameters(), lr=0.05)
    g = torch.Generator(device="cpu").manual_seed(11 + rank)
    for it in range(6):
        x = torch.randn(32, 64, generator=g).to(dev); y = torch.randn(32, 10, generator=g).to(dev)
        for m, o in ((mine, o1), (ref, o2)):
            o.zero_grad()
            torch.nn.functional.mse_loss(m(x), y).backward()
            o.step()
    mine.check_comm_status()
    tol = 5e-2 if comm_dtype == "bf16" else 1e-5
    for (n, p), q in zip(mine.module.named_parameters(), ref.module.parameters()):
        assert torch.allclose(p, q, atol=tol, rtol=tol), (n, (p - q).abs().max().item())
    for (n, b), c in zip(mine.module.named_buffers(), ref.module.buffers()):
        assert torch.allclose(b.float(), c.float(), atol=tol, rtol=tol), n
    info = mine.get_logging_data()
    assert info["comm"] == "b200dt-peer-memory", info
    assert info["rebuilt"] and len(info["bucket_sizes_mib"]) > 1
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write(str(info["bucket_modes"]))


@needs2
@pytest.mark.parametrize("mode,comm_dtype", [("auto", "fp32"), ("oneshot", "fp32"), ("twoshot", "fp32"), ("auto", "bf16")])
def test_ddp_matches_torch_ddp(tmp_path, mode, comm_dtype):
    w = _world()
    spawn(_ddp_vs_torch, args=(w, group.find_free_port(), str(tmp_path), mode, comm_dtype), nprocs=w)
    assert all((tmp_path / f"ok{r}").exists() for r in range(w))
