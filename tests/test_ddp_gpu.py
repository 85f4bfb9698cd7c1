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
    arena = coll.make_arena(64 << 20, dev)
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
    # tensors living in the arena: in-place two-shot, resident and stray tensors mixed
    for rep in range(3):
        numels = [3, 1000, 70001, 1_000_000]
        table = coll.TensorTable(numels, dev, False, False)
        views = coll.stage_views(arena, table, [(n,) for n in numels], stage_base=1 << 20)
        stray = torch.randn(numels[1], device=dev)                 # not resident: copied in, left untouched
        srcs = [torch.randn(n, device=dev) for n in numels]
        srcs[1] = stray.clone()
        for v, s_ in zip(views, srcs):
            v.copy_(s_)
        ref = [t.clone() for t in srcs]
        for r in ref:
            dist.all_reduce(r)
            r.div_(world)
        bound = [views[0], stray, views[2], views[3]]
        table.bind(bound, torch.cuda.current_stream(dev))
        mode = "nvls" if arena.has_multicast else "twoshot"
        coll.launch(mode, arena, table, 1.0 / world, torch.cuda.current_stream(dev), stage_base=1 << 20, in_place=True)
        torch.cuda.synchronize(dev)
        arena.check_status()
        for v, b in zip(views, ref):                               # results are read from the views
            assert (v - b).abs().max().item() <= 1e-5 * max(1.0, b.abs().max().item())
        assert torch.equal(stray, srcs[1])
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
    # where the driver offers multicast the library's own arena must have bound it, i.e. the multimem
    # (NVLS) kernels really ran above instead of being skipped
    if _multicast_supported():
        assert "nvls" in (tmp_path / "ok0").read_text()


def _multicast_supported():
    from pytorch_distributed_training_tutorials_b200._lib import lib
    h = lib()._handle
    h.b200dt_multicast_supported.argtypes = [__import__("ctypes").c_int]
    return int(h.b200dt_multicast_supported(0)) == 1


def _ddp_vs_torch(rank, world, port, out_dir, mode, comm_dtype, views=False):
    from pytorch_distributed_training_tutorials_b200.parallel.ddp import DistributedDataParallel as DDP
    st = group.init(backend="nccl", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    dev = st.device
    torch.manual_seed(5)

    def make():
        torch.manual_seed(5)
        return torch.nn.Sequential(torch.nn.Linear(64, 512), torch.nn.ReLU(), torch.nn.Linear(512, 2048),
                                   torch.nn.ReLU(), torch.nn.BatchNorm1d(2048), torch.nn.Linear(2048, 10)).to(dev)
    try:
        mine = DDP(make(), device_ids=[dev.index], allreduce=mode, comm_dtype=comm_dtype, bucket_cap_mb=1,
                   gradient_as_bucket_view=views)
    except RuntimeError as e:
        if mode == "nvls" and "no multicast" in str(e):
            group.destroy()
            open(os.path.join(out_dir, f"ok{rank}"), "w").write("nomc")
            return
        raise
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
    if views:
        # every gradient is a view of the symmetric arena, at the slot its bucket assigns
        lo = mine._arena.buf.data_ptr(); hi = lo + mine._arena.nbytes
        assert all(lo <= p.grad.data_ptr() < hi for p in mine.module.parameters())
        assert all(m == "twoshot" or m == "nvls" for m in info["bucket_modes"])
        # gradient accumulation under no_sync keeps adding into the views
        x = torch.randn(32, 64, generator=g).to(dev); y = torch.randn(32, 10, generator=g).to(dev)
        for m, o in ((mine, o1), (ref, o2)):
            o.zero_grad()
            with m.no_sync():
                torch.nn.functional.mse_loss(m(x), y).backward()
            torch.nn.functional.mse_loss(m(x), y).backward()
            o.step()
        for (n, p), q in zip(mine.module.named_parameters(), ref.module.parameters()):
            assert torch.allclose(p, q, atol=tol, rtol=tol), (n, (p - q).abs().max().item())
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write(str(info["bucket_modes"]))


@needs2
@pytest.mark.parametrize("mode,comm_dtype", [("auto", "fp32"), ("oneshot", "fp32"), ("twoshot", "fp32"), ("auto", "bf16"),
                                             ("nvls", "fp32"), ("nvls", "bf16")])
def test_ddp_matches_torch_ddp(tmp_path, mode, comm_dtype):
    w = _world()
    spawn(_ddp_vs_torch, args=(w, group.find_free_port(), str(tmp_path), mode, comm_dtype), nprocs=w)
    assert all((tmp_path / f"ok{r}").exists() for r in range(w))
    if mode == "nvls":
        ran = (tmp_path / "ok0").read_text() != "nomc"
        assert ran or not _multicast_supported(), "the driver offers multicast but the arena did not bind it"
        if not ran:
            pytest.skip("no NVLS multicast on this box")


def _ddp_views_native(rank, world, port, out_dir):
    """Own tcgen05 Linear layers: the weight-gradient GEMM writes straight into the arena slot."""
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.models.toy import MLP
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    st = group.init(backend="nccl", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    dev = st.device
    models, opts = [], []
    for views in (False, True):
        torch.manual_seed(3)
        m = b2.DistributedDataParallel(MLP(64, 256, 3, 16).to(dev), device_ids=[dev.index], bucket_cap_mb=0.25,
                                       gradient_as_bucket_view=views)
        models.append(m); opts.append(b2.FusedSGD(m.parameters(), lr=0.05))
    g = torch.Generator(device="cpu").manual_seed(21 + rank)
    for it in range(5):
        x = torch.randn(64, 64, generator=g).to(dev); y = torch.randn(64, 16, generator=g).to(dev)
        for m, o in zip(models, opts):
            o.zero_grad()
            Fn.mse_loss(m(x), y).backward()
            o.step()
    lo = models[1]._arena.buf.data_ptr(); hi = lo + models[1]._arena.nbytes
    for (n, p), q in zip(models[1].module.named_parameters(), models[0].module.parameters()):
        assert lo <= p.grad.data_ptr() < hi, n
        assert torch.allclose(p, q, atol=1e-6, rtol=1e-5), (n, (p - q).abs().max().item())
    # the 2-D weights' gradients were produced in place: after zero_grad + backward (no finalize yet
    # needed to see it) their storage is the arena slot already
    opts[1].zero_grad()
    with models[1].no_sync():
        Fn.mse_loss(models[1](x), y).backward()
    w = models[1].module.layers[0].weight
    assert lo <= w.grad.data_ptr() < hi
    models[1].check_comm_status()
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


@needs2
def test_ddp_gradient_views_with_native_layers(tmp_path):
    w = _world()
    spawn(_ddp_views_native, args=(w, group.find_free_port(), str(tmp_path)), nprocs=w)
    assert all((tmp_path / f"ok{r}").exists() for r in range(w))


@needs2
def test_ddp_gradient_as_bucket_view(tmp_path):
    w = _world()
    spawn(_ddp_vs_torch, args=(w, group.find_free_port(), str(tmp_path), "auto", "fp32", True), nprocs=w)
    assert all((tmp_path / f"ok{r}").exists() for r in range(w))


def _timeout_child(rank, world, port, out_dir):
    import json
    import time
    from pytorch_distributed_training_tutorials_b200._lib import lib
    from pytorch_distributed_training_tutorials_b200.ops import collectives as coll
    st = group.init(backend="nccl", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    dev = st.device
    arena = coll.make_arena(1 << 20, dev)
    lib()._handle.b200dt_set_coll_timeout_ms.argtypes = [__import__("ctypes").c_longlong]
    lib()._handle.b200dt_set_coll_timeout_ms(1500)
    x = torch.ones(1000, device=dev)
    coll.all_reduce_([x], arena)                    # a healthy collective first
    torch.cuda.synchronize(dev)
    assert float(x[0]) == world
    dist.barrier(device_ids=[dev.index])
    if rank == world - 1:
        os._exit(0)                                 # this rank "dies" before the next collective
    t0 = time.time()
    coll.all_reduce_([x], arena)
    raised = False
    try:
        torch.cuda.synchronize(dev)
    except RuntimeError:
        raised = True                               # the kernel trapped: nothing was reduced from partial data
    msg = None
    try:
        arena.check_status()
    except coll.CollectiveTimeout as e:
        msg = str(e)
    with open(os.path.join(out_dir, f"t{rank}.json"), "w") as f:
        json.dump({"raised": raised, "msg": msg, "elapsed": time.time() - t0}, f)
        f.flush()
        os.fsync(f.fileno())
    os._exit(0)                                     # the CUDA context is gone: skip every destructor


@needs2
def test_dead_peer_is_detected_within_the_time_limit(tmp_path):
    """One rank exits between two collectives: the survivors' kernels give up after the configured limit, trap
    (no gradient is ever built from partial data), and the host reads who was missing from the status words."""
    import json
    w = _world()
    spawn(_timeout_child, args=(w, group.find_free_port(), str(tmp_path)), nprocs=w)
    for r in range(w - 1):
        rec = json.loads((tmp_path / f"t{r}.json").read_text())
        assert rec["raised"], rec
        assert rec["msg"] is not None and "waiting for rank" in rec["msg"], rec
        assert rec["elapsed"] < 20, rec
