"""CPU-only unit tests: sampler equality with torch, launch contracts, scatter/gather pytrees,
checkpoint round trips, placement, pipeline equivalence, bucket assignment."""
import collections
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
from torch.utils.data import DistributedSampler as TorchSampler

import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.data import DistributedSampler, MyTrainDataset, RandomDataset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n", [1, 7, 10, 2048, 2049])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("shuffle,drop_last,seed", [(True, False, 0), (False, False, 0), (True, True, 5), (False, True, 1)])
def test_sampler_matches_torch(n, world, shuffle, drop_last, seed):
    ds = list(range(n))
    for rank in range(world):
        a = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        b = TorchSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=seed, drop_last=drop_last)
        for epoch in (0, 1, 7):
            a.set_epoch(epoch); b.set_epoch(epoch)
            assert list(a) == list(b)
            assert len(a) == len(b)


def test_sampler_needs_group_or_explicit_rank():
    with pytest.raises(RuntimeError):
        DistributedSampler(list(range(4)))
    with pytest.raises(ValueError):
        DistributedSampler(list(range(4)), num_replicas=2, rank=2)


def test_steps_per_epoch_match_reference_banner():
    # 2048 samples, per-rank batch 32: 16 steps on 4 ranks, 64 on one (02.ddp_toy_example.ipynb:231,268)
    ds = MyTrainDataset(2048)
    for world, steps in ((4, 16), (1, 64), (2, 32), (8, 8)):
        s = DistributedSampler(ds, num_replicas=world, rank=0)
        dl = torch.utils.data.DataLoader(ds, batch_size=32, sampler=s)
        assert len(dl) == steps
    x, y = ds[0]
    assert x.shape == (20,) and y.shape == (1,)
    assert len(RandomDataset(32, 1024)) == 1024 and RandomDataset(32, 1024)[0].shape == (32,)


def test_scatter_gather_pytrees():
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import gather, scatter, scatter_kwargs
    devs = ["cpu"] * 4
    x = torch.arange(32 * 3).float().view(32, 3)
    chunks = scatter(x, devs)
    assert [c.shape[0] for c in chunks] == [8, 8, 8, 8]
    assert [c.shape[0] for c in scatter(torch.zeros(10, 2), devs)] == [3, 3, 3, 1]
    assert len(scatter(torch.zeros(2, 2), devs)) == 2                  # fewer chunks than devices
    NT = collections.namedtuple("NT", ["a", "b"])
    tree = (x, [x, {"k": x}], NT(x, 5), "s")
    parts = scatter(tree, devs)
    assert len(parts) == 4 and parts[1][0].shape == (8, 3) and parts[2][1][1]["k"].shape == (8, 3)
    assert isinstance(parts[0][2], NT) and parts[3][2].b == 5 and parts[0][3] == "s"
    ins, kws = scatter_kwargs((x,), {"m": x}, devs)
    assert len(ins) == len(kws) == 4 and kws[2]["m"].shape == (8, 3)
    ins, kws = scatter_kwargs((), {"m": x}, devs)
    assert ins == ((), (), (), ())
    outs = [{"o": c * 2, "n": None, "t": (c, c)} for c in chunks]
    g = gather(outs, "cpu")
    assert torch.equal(g["o"], x * 2) and g["n"] is None and torch.equal(g["t"][1], x)
    with pytest.warns(UserWarning):
        v = gather([torch.tensor(1.0), torch.tensor(2.0)], "cpu")
    assert v.tolist() == [1.0, 2.0]


def test_data_parallel_virtual_devices_matches_chunked_reference(capsys):
    import copy
    from pytorch_distributed_training_tutorials_b200.models.toy import SampleModel
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import DataParallel
    torch.manual_seed(0)
    m = SampleModel(32, 2)
    ref = copy.deepcopy(m); ref.verbose = False
    dp = DataParallel(m, device_ids=["cpu"] * 4)
    x = torch.randn(32, 32)
    out = dp(x)
    printed = capsys.readouterr().out
    assert printed.count("In Model: input size torch.Size([8, 32]) output size torch.Size([8, 2])") == 4
    assert out.shape == (32, 2)
    out.sum().backward()
    ref(x).sum().backward()
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5)
    assert list(dp.state_dict().keys()) == ["module.fc.weight", "module.fc.bias"]
    s, o = dp.forward_sum(x)
    assert torch.allclose(s, ref(x).sum(), atol=1e-4) and o.shape == (32, 2)
    # errors inside a replica surface in the caller with the replica index
    class Boom(torch.nn.Module):
        def forward(self, x):
            raise ValueError("bad replica")
    with pytest.raises(ValueError, match="replica"):
        DataParallel(Boom(), device_ids=["cpu", "cpu"])(x)
    # in-order issue from the calling thread (the CUDA default) gives the same result and error surface
    m.verbose = False
    dp2 = DataParallel(m, device_ids=["cpu"] * 4, threaded=False)
    assert torch.allclose(dp2(x), ref(x), atol=1e-6)
    with pytest.raises(ValueError, match="replica 0"):
        DataParallel(Boom(), device_ids=["cpu", "cpu"], threaded=False)(x)


def test_checkpoint_roundtrip_with_plain_torch(tmp_path):
    from pytorch_distributed_training_tutorials_b200.checkpoint import load_checkpoint, save_checkpoint
    from pytorch_distributed_training_tutorials_b200.parallel.data_parallel import DataParallel
    m = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    m(torch.randn(5, 4)).sum().backward(); opt.step()
    wrapped = DataParallel(m, device_ids=["cpu", "cpu"])
    path = str(tmp_path / "ckpt.pt")
    assert save_checkpoint(wrapped, opt, epoch=4, path=path, extra={"note": "x"}) == path
    blob = torch.load(path, weights_only=False)
    assert not any(k.startswith("module.") for k in blob["model"])
    plain = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    plain.load_state_dict(blob["model"])                              # plain torch reads our file
    m2 = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    o2 = torch.optim.Adam(m2.parameters(), lr=1e-3)
    info = load_checkpoint(path, DataParallel(m2, device_ids=["cpu"]), o2)
    assert info["epoch"] == 4 and info["extra"] == {"note": "x"} and not info["missing"]
    assert torch.equal(m2[0].weight, m[0].weight)
    assert o2.state_dict()["state"][0]["exp_avg"].equal(opt.state_dict()["state"][0]["exp_avg"])
    # a file written by plain torch from a wrapped model (module. prefix) loads too
    torch.save({"module." + k: v for k, v in m.state_dict().items()}, path)
    assert load_checkpoint(path, m2)["missing"] == []


def test_checkpoint_dedups_aliased_registration(tmp_path):
    from pytorch_distributed_training_tutorials_b200.checkpoint import load_checkpoint, model_state

    class Twice(torch.nn.Module):      # the reference's split ResNet registers layers twice
        def __init__(self):
            super().__init__()
            self.layer1 = torch.nn.Linear(3, 3)
            self.seq1 = torch.nn.Sequential(self.layer1)
    t = Twice()
    assert set(t.state_dict()) == {"layer1.weight", "layer1.bias", "seq1.0.weight", "seq1.0.bias"}
    sd = model_state(t)
    assert set(sd) == {"layer1.weight", "layer1.bias"}
    torch.save(sd, tmp_path / "a.pt")
    t2 = Twice()
    assert load_checkpoint(str(tmp_path / "a.pt"), t2)["missing"] == []
    assert torch.equal(t2.seq1[0].weight, t.layer1.weight)


def test_auto_place_reproduces_reference_listing():
    from pytorch_distributed_training_tutorials_b200.models.llama_like import LlamaLikeForCausalLM
    from pytorch_distributed_training_tutorials_b200.parallel.place import auto_place, module_bytes, placement_table
    m = LlamaLikeForCausalLM(vocab=64, dim=16, layers=32, heads=2, hidden=32)
    assert len(list(m.parameters())) == 291                           # 03.model_parallel.ipynb:114-404
    blocks = m.placement_blocks()
    layer = module_bytes(m.model.layers[0]); emb = module_bytes(m.model.embed_tokens)
    devs = [torch.device("cuda", i) for i in range(4)]                 # placement is pure bookkeeping
    budgets = {devs[0]: emb + 6 * layer, devs[1]: 8 * layer, devs[2]: 8 * layer, devs[3]: 10 ** 12}
    pl = auto_place(m, devs, max_memory=budgets, blocks=blocks)
    per_dev = collections.Counter()
    for name, dev in pl.items():
        sub = dict(m.named_modules())[name]
        per_dev[dev.index] += len(list(sub.parameters()))
    assert [per_dev[i] for i in range(4)] == [55, 72, 72, 92]
    bal = auto_place(m, devs, blocks=blocks)                           # balanced default keeps order
    idx = [bal[n].index for n, _ in blocks]
    assert idx == sorted(idx) and set(idx) == {0, 1, 2, 3}
    assert len(placement_table(m)) == 291


def test_staged_and_pipelined_equal_plain_module():
    from pytorch_distributed_training_tutorials_b200.parallel.model_parallel import Pipeline, StagedModule, split_sequential, model_size
    torch.manual_seed(0)
    seq = torch.nn.Sequential(torch.nn.Linear(6, 8), torch.nn.ReLU(), torch.nn.Linear(8, 8), torch.nn.ReLU(), torch.nn.Linear(8, 2))
    stages = split_sequential(seq, [2, 4])
    assert len(stages) == 3
    x = torch.randn(10, 6)
    ref = seq(x)
    st = StagedModule(stages, ["cpu"] * 3)
    assert torch.allclose(st(x), ref)
    pp = Pipeline(stages, ["cpu"] * 3, split_size=3)
    out = pp(x)
    assert torch.allclose(out, ref, atol=1e-6)
    out.sum().backward()
    g = [p.grad.clone() for p in seq.parameters()]
    seq.zero_grad(); seq(x).sum().backward()
    assert all(torch.allclose(a, p.grad, atol=1e-5) for a, p in zip(g, seq.parameters()))
    assert model_size(st) == model_size(seq)


def test_split_resnet_has_reference_param_count():
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.parallel.model_parallel import model_size
    assert model_size(resnet50(num_classes=1000)) == 25557032          # 03.model_parallel.ipynb:857,888


def test_block_level_units_and_time_balanced_cuts():
    """Pipeline cuts may fall between any two bottleneck blocks; the cut finder minimises the slowest stage."""
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50
    from pytorch_distributed_training_tutorials_b200.parallel.place import balanced_cuts
    m = resnet50(num_classes=10)
    names = m.unit_names()
    assert names[0] == "stem" and names[-1] == "pool" and len(names) == 18
    assert m.unit_index("layer2") == names.index("layer2.3") and m.unit_index("layer2", last=False) == names.index("layer2.0")
    with pytest.raises(ValueError):
        m.unit_index("layer9")
    # every parameter belongs to exactly one unit (fc travels with the pool)
    owned = sum(sum(p.numel() for mod in m.unit_modules(n) for p in mod.parameters()) for n in names)
    assert owned + sum(p.numel() for p in m.fc.parameters()) == sum(p.numel() for p in m.parameters())
    assert balanced_cuts([5, 1, 1, 1, 1, 1], 2) == [0]
    assert balanced_cuts([1, 1, 1, 1, 1, 5], 2) == [4]
    assert balanced_cuts([1, 2, 3, 4, 5, 6, 7], 3) == [3, 5]
    assert balanced_cuts([3.0], 1) == []
    with pytest.raises(ValueError):
        balanced_cuts([1, 2], 3)
    # the split model runs unit-exactly like the unsplit one (both devices the CPU here)
    torch.manual_seed(0)
    sp = ModelParallelResNet50(num_classes=10, dev0="cpu", dev1="cpu", split_after="layer2.1")
    assert sp._cut == names.index("layer2.1")
    with pytest.raises(ValueError):
        ModelParallelResNet50(num_classes=10, dev0="cpu", dev1="cpu", split_after="pool")


def test_bucket_assignment_matches_torch_resnet50():
    import torchvision
    from pytorch_distributed_training_tutorials_b200.parallel._native_reducer import load
    R = load()
    ps = list(torchvision.models.resnet50().parameters())
    rev = list(reversed(ps)); idx = list(reversed(range(len(ps))))
    mine, _ = R.assign_buckets(rev, [1 << 20, 25 << 20], idx)
    ref, _ = dist._compute_bucket_assignment_by_size(rev, [1 << 20, 25 << 20], [False] * len(ps), idx)
    assert mine == ref and len(mine) == 5
    a, _ = R.assign_buckets(ps, [2 ** 62])
    assert a == [list(range(len(ps)))]


def test_run_launcher_env_contract(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(
        "import os, json\n"
        "keys = ['LOCAL_RANK','RANK','GROUP_RANK','ROLE_RANK','LOCAL_WORLD_SIZE','WORLD_SIZE','MASTER_ADDR',"
        "'MASTER_PORT','TORCHELASTIC_RESTART_COUNT','TORCHELASTIC_MAX_RESTARTS','TORCHELASTIC_RUN_ID','OMP_NUM_THREADS']\n"
        f"open(r'{tmp_path}' + '/env' + os.environ['RANK'] + '.json','w').write(json.dumps({{k: os.environ.get(k) for k in keys}}))\n")
    env = dict(os.environ, PYTHONPATH=ROOT)
    env.pop("OMP_NUM_THREADS", None)
    r = subprocess.run([sys.executable, "-m", "pytorch_distributed_training_tutorials_b200.run", "--standalone",
                        "--nproc-per-node", "3", str(script)], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "Setting OMP_NUM_THREADS" in r.stderr
    import json
    envs = [json.load(open(tmp_path / f"env{i}.json")) for i in range(3)]
    assert [e["LOCAL_RANK"] for e in envs] == ["0", "1", "2"] and all(e["WORLD_SIZE"] == "3" for e in envs)
    assert all(e["MASTER_ADDR"] == "127.0.0.1" and e["OMP_NUM_THREADS"] == "1" for e in envs)
    assert len({e["MASTER_PORT"] for e in envs}) == 1
    # default is ONE worker, like torchrun without --nproc-per-node (02.ddp_toy_example.ipynb:268)
    for f in tmp_path.glob("env*.json"):
        f.unlink()
    r = subprocess.run([sys.executable, "-m", "pytorch_distributed_training_tutorials_b200.run", "--standalone",
                        str(script)], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and sorted(p.name for p in tmp_path.glob("env*.json")) == ["env0.json"]


def test_run_launcher_failure_and_restart(tmp_path):
    script = tmp_path / "f.py"
    script.write_text("import os, sys, time\n"
                      "if os.environ['RANK'] == '1' and os.environ['TORCHELASTIC_RESTART_COUNT'] == '0': sys.exit(3)\n"
                      "time.sleep(0.2)\n")
    env = dict(os.environ, PYTHONPATH=ROOT)
    base = [sys.executable, "-m", "pytorch_distributed_training_tutorials_b200.run", "--standalone", "--nproc-per-node", "2"]
    r = subprocess.run(base + [str(script)], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 3 and "failed with exit code 3" in r.stderr
    r = subprocess.run(base + ["--max-restarts", "1", str(script)], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "restarting worker group (1/1)" in r.stderr


# --------------------------------------------------------------------------- prefetcher / run_batches
def test_device_prefetcher_cpu_passthrough_and_len():
    from pytorch_distributed_training_tutorials_b200.data import DevicePrefetcher
    batches = [(torch.full((2, 3), float(i)), torch.tensor([i, i])) for i in range(5)]
    pf = DevicePrefetcher(batches, "cpu")
    assert len(pf) == 5
    got = list(pf)
    assert len(got) == 5
    for i, (x, y) in enumerate(got):
        assert torch.equal(x, batches[i][0]) and torch.equal(y, batches[i][1])
    assert list(DevicePrefetcher([], "cpu")) == []


def test_trainer_run_batches_cpu_matches_manual_loop():
    import copy
    from pytorch_distributed_training_tutorials_b200.trainer import Trainer
    torch.manual_seed(0)
    m = torch.nn.Linear(20, 4)
    r = copy.deepcopy(m)
    batches = [(torch.randn(8, 20), torch.randint(0, 4, (8,))) for _ in range(4)]
    tr = Trainer(m, None, torch.optim.SGD(m.parameters(), lr=0.1), gpu_id=0, wrap_ddp=False)
    losses = tr.run_batches(iter(batches))
    opt = torch.optim.SGD(r.parameters(), lr=0.1)
    ref = []
    for x, y in batches:
        opt.zero_grad(); l = torch.nn.functional.cross_entropy(r(x), y); l.backward(); opt.step(); ref.append(float(l))
    assert losses == pytest.approx(ref, rel=1e-6)
    assert tr.global_step == 4
    for p, q in zip(m.parameters(), r.parameters()):
        assert torch.allclose(p, q)


# --------------------------------------------------------------------------- multi-node emulation, busy port
def test_two_launcher_agents_emulate_two_nodes(tmp_path):
    """``--nnodes 2`` with two agents on one host (different --node-rank, shared master port): ranks
    0 and 1 meet over gloo and all-reduce."""
    from pytorch_distributed_training_tutorials_b200 import group
    script = tmp_path / "n.py"
    script.write_text(
        "import os, torch, torch.distributed as dist\n"
        "import pytorch_distributed_training_tutorials_b200 as b2\n"
        "st = b2.init(backend='gloo')\n"
        "t = torch.tensor([float(st.rank + 1)]); dist.all_reduce(t)\n"
        f"open(r'{tmp_path}' + '/r' + str(st.rank), 'w').write(f\"{{st.world_size}} {{os.environ['GROUP_RANK']}} {{os.environ['LOCAL_RANK']}} {{t.item()}}\")\n"
        "b2.destroy()\n")
    port = group.find_free_port()
    env = dict(os.environ, PYTHONPATH=ROOT)
    base = [sys.executable, "-m", "pytorch_distributed_training_tutorials_b200.run", "--nnodes", "2", "--nproc-per-node", "1",
            "--master-addr", "127.0.0.1", "--master-port", str(port)]
    procs = [subprocess.Popen(base + ["--node-rank", str(nr), str(script)], env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True) for nr in (0, 1)]
    for p in procs:
        out, err = p.communicate(timeout=180)
        assert p.returncode == 0, err
    assert (tmp_path / "r0").read_text() == "2 0 0 3.0"
    assert (tmp_path / "r1").read_text() == "2 1 0 3.0"


def _busy_port_child(rank, port):
    from pytorch_distributed_training_tutorials_b200 import group
    group.init(backend="gloo", rank=0, world_size=1, master_addr="127.0.0.1", master_port=port, timeout_s=20)


def test_busy_master_port_is_an_error_not_a_hang():
    """The reference hard-codes port 12345, so a stale run makes the next one collide
    (/root/reference/ddp_gpus.py:14); here the collision surfaces as an exception naming the problem,
    and ``find_free_port`` / the launcher's default avoid it in the first place."""
    import socket
    from pytorch_distributed_training_tutorials_b200 import group
    from pytorch_distributed_training_tutorials_b200.launch.spawn import spawn, ProcessRaisedException
    s = socket.socket(); s.bind(("127.0.0.1", 0)); s.listen(1)
    port = s.getsockname()[1]
    try:
        with pytest.raises(ProcessRaisedException) as ei:
            spawn(_busy_port_child, args=(port,), nprocs=1)
        msg = str(ei.value).lower()
        assert "address already in use" in msg or "eaddrinuse" in msg or "in use" in msg, msg
    finally:
        s.close()
    p1, p2 = group.find_free_port(), group.find_free_port()
    assert 1024 < p1 < 65536 and 1024 < p2 < 65536


# --------------------------------------------------------------------------- small pure-Python helpers
def test_two_level_slices_cover_the_range_exactly():
    from pytorch_distributed_training_tutorials_b200.parallel.comm import _slices
    for total in (0, 1, 63, 64, 65, 1000, 262144, 67_000_003):
        for parts in (1, 2, 3, 7, 8):
            sl = _slices(total, parts)
            assert len(sl) == parts
            covered = 0
            for lo, hi in sl:
                assert lo == covered or (lo == total and hi == total)
                assert lo % 64 == 0 or lo == total
                covered = max(covered, hi)
            assert covered == total


def test_tensor_table_stage_layout_is_contiguous_per_tensor():
    """Every tensor of a bucket occupies one contiguous, 32-byte aligned range of the staging region:
    that is what lets DDP hand out gradient VIEWS of the arena (in-place all-reduce)."""
    from pytorch_distributed_training_tutorials_b200.ops.collectives import TensorTable
    numels = [3, 1000, 70001, 16384, 1, 2_000_003]
    for blocks in (4, 32, 128):
        t = TensorTable(numels, torch.device("cpu"), False, False, target_blocks=blocks)
        assert len(t.tensor_offsets) == len(numels)
        end = 0
        for off, n in zip(t.tensor_offsets, numels):
            assert off % 8 == 0 and off >= end
            end = off + n
        assert t.stage_elems >= end and t.stage_bytes == t.stage_elems * 4
        raw = t.chunks.view(-1, 3)
        # chunk rows: (start << 32 | tensor, len, stage_off); a tensor's chunks follow each other in the stage
        per_tensor = {}
        for row in raw.tolist():
            ti, start = row[0] & 0xFFFFFFFF, row[0] >> 32
            per_tensor.setdefault(ti, []).append((start, row[1] & 0xFFFFFFFF, row[2]))
        for ti, rows in per_tensor.items():
            base = t.tensor_offsets[ti]
            for start, ln, off in rows:
                assert off == base + start
            assert sum(ln for _, ln, _ in rows) == numels[ti]


def test_fault_spec_parsing():
    from pytorch_distributed_training_tutorials_b200.trainer import _parse_fault
    assert _parse_fault("") is None
    assert _parse_fault("rank=1,step=3") == (1, 3)
    assert _parse_fault(" step=0 , rank=7 ") == (7, 0)
    with pytest.raises(ValueError):
        _parse_fault("rank=1")
    with pytest.raises(ValueError):
        _parse_fault("rank=a,step=2")


def test_graphed_wrappers_refuse_cpu_devices():
    import pytorch_distributed_training_tutorials_b200 as b2
    m = torch.nn.Linear(8, 2)
    x, y = torch.randn(4, 8), torch.randn(4, 2)
    with pytest.raises(ValueError, match="CUDA"):
        b2.GraphedDataParallel(m, ["cpu", "cpu"], torch.nn.functional.mse_loss, x, y)
    with pytest.raises(ValueError, match="CUDA"):
        b2.GraphedPipeline([m], ["cpu"], 2, torch.nn.functional.mse_loss, x, y)


def test_example_saves_and_resumes_under_the_launcher(tmp_path):
    """examples/ddp_run.py on 2 gloo ranks: one epoch + checkpoint, then --resume continues at epoch 1."""
    env = dict(os.environ, PYTHONPATH=ROOT)
    ck = str(tmp_path / "ck.pt")
    base = [sys.executable, "-m", "pytorch_distributed_training_tutorials_b200.run", "--standalone", "--nproc-per-node", "2",
            os.path.join(ROOT, "examples", "ddp_run.py"), "--backend", "gloo", "--batch_size", "64"]
    r = subprocess.run(base + ["--max_epochs", "1", "--save", ck], env=env, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr
    assert "Epoch: 0, Batch size: 64 | Steps 16]" in r.stdout and os.path.exists(ck)
    sd = torch.load(ck, map_location="cpu")
    assert sd["epoch"] == 0 and set(sd["model"].keys()) == {"weight", "bias"}
    r = subprocess.run(base + ["--max_epochs", "3", "--resume", ck], env=env, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr
    assert "continuing with epoch 1" in r.stdout
    assert "Epoch: 0," not in r.stdout and "Epoch: 1," in r.stdout and "Epoch: 2," in r.stdout


# --------------------------------------------------------------------------- skinny Linear: host emulation
@pytest.mark.parametrize("B,K,N", [(32, 20, 1), (8, 32, 2), (20, 10000, 10), (20, 10, 5), (1, 1, 1), (7, 3, 9),
                                   (300, 17, 33), (2048, 20, 1)])
def test_skinny_linear_element_code_matches_torch_on_the_host(B, K, N):
    """csrc/skinny.cu compiles its per-element functions for host and device; the *_host entry points run
    them serially, so forward / dgrad / wgrad / bias-grad indexing is checked here without a GPU."""
    from pytorch_distributed_training_tutorials_b200 import _lib
    L = _lib.lib()
    torch.manual_seed(B * 7 + K)
    x = torch.randn(B, K); w = torch.randn(N, K) / K ** 0.5; b = torch.randn(N); dy = torch.randn(B, N)
    y = torch.full((B, N), float("nan"))
    assert L.b200dt_skinny_linear_fwd_host(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), B, K, N) == 0
    ref = torch.nn.functional.linear(x, w, b)
    assert torch.allclose(y, ref, rtol=1e-4, atol=1e-4), (y - ref).abs().max()
    y2 = torch.full((B, N), float("nan"))
    assert L.b200dt_skinny_linear_fwd_host(x.data_ptr(), w.data_ptr(), None, y2.data_ptr(), B, K, N) == 0
    assert torch.allclose(y2, torch.nn.functional.linear(x, w), rtol=1e-4, atol=1e-4)
    dx = torch.full((B, K), float("nan")); dw = torch.full((N, K), float("nan")); db = torch.full((N,), float("nan"))
    assert L.b200dt_skinny_linear_bwd_host(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(),
                                           db.data_ptr(), B, K, N) == 0
    assert torch.allclose(dx, dy @ w, rtol=1e-4, atol=1e-4)
    assert torch.allclose(dw, dy.t() @ x, rtol=1e-4, atol=1e-3)
    assert torch.allclose(db, dy.sum(0), rtol=1e-4, atol=1e-3)
    # any subset of the three gradients
    dw2 = torch.full((N, K), float("nan"))
    assert L.b200dt_skinny_linear_bwd_host(dy.data_ptr(), x.data_ptr(), w.data_ptr(), None, dw2.data_ptr(), None, B, K, N) == 0
    assert torch.equal(dw2, dw)
    assert L.b200dt_skinny_linear_fwd_host(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), -1, K, N) == -3


def test_launcher_accepts_torchrun_rendezvous_spelling():
    from pytorch_distributed_training_tutorials_b200.run import parse_args
    a = parse_args(["--nproc_per_node", "2", "--rdzv-backend", "c10d", "--rdzv-endpoint", "10.0.0.5:29611",
                    "--rdzv-id", "job7", "train.py", "--max_epochs", "3"])
    assert (a.master_addr, a.master_port, a.run_id, a.nproc_per_node) == ("10.0.0.5", 29611, "job7", "2")
    assert a.training_script == "train.py" and a.training_script_args == ["--max_epochs", "3"]
    b = parse_args(["--rdzv_endpoint", "localhost", "train.py"])
    assert (b.master_addr, b.master_port) == ("localhost", 29500)


def test_native_linear_extension_is_built_and_switchable(monkeypatch):
    """The C++ autograd Linear lives in the reducer extension; its kernel addresses come from the ctypes handle, and
    ``B200DT_CPP_LINEAR=0`` keeps the Python node (the two are interchangeable; parity is tested on the GPU)."""
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    from pytorch_distributed_training_tutorials_b200.parallel._native_reducer import load
    mod = load()
    assert hasattr(mod, "linear") and hasattr(mod, "linear_set_fns") and mod.linear_launches() >= 0
    monkeypatch.setattr(Fn, "_cpp_linear", None)
    monkeypatch.setenv("B200DT_CPP_LINEAR", "0")
    assert Fn._native_linear() is False
    monkeypatch.setattr(Fn, "_cpp_linear", None)
    monkeypatch.setenv("B200DT_CPP_LINEAR", "1")
    assert Fn._native_linear() is mod


def test_model_parallel_demo_resnet_walks_through_on_cpu():
    """examples/model_parallel_demo.py resnet --cpu: the reference's split (03.model_parallel.ipynb:807-833), naive and
    pipelined, on CPU placeholders; both print torchvision's parameter count (ibid. 857, 888)."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "examples", "model_parallel_demo.py"), "resnet", "--cpu"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.count("25557032") == 2
    assert "naive: (8, 1000) pipelined (cut after layer2.1): (8, 1000)" in out.stdout


def test_statistics_copy_count_and_switches_through_the_c_abi():
    """Host-side policy of the BatchNorm partial sums (csrc/config.cu), reachable without a GPU: the copy count
    falls with the channel count, deterministic mode uses one slot per CTA, the pre-zeroed switch round-trips."""
    from pytorch_distributed_training_tutorials_b200 import _lib
    L = _lib.lib()
    was = _lib.is_deterministic()
    try:
        _lib.set_deterministic(False)
        assert [L.b200dt_bn_replicas_used(c) for c in (8, 64, 128, 256, 512, 1024, 2048, 4096)] == [16, 16, 16, 8, 4, 2, 1, 1]
        assert L.b200dt_bn_replicas() == 16
        _lib.set_deterministic(True)
        assert L.b200dt_bn_replicas() == 160 and L.b200dt_bn_replicas_used(2048) == 160
    finally:
        _lib.set_deterministic(was)
    prev = L.b200dt_get_sums_prezeroed()
    L.b200dt_set_sums_prezeroed(1)
    assert L.b200dt_get_sums_prezeroed() == 1
    L.b200dt_set_sums_prezeroed(prev)
    assert L.b200dt_capture_id(None) == 0               # no capture (and no GPU here): 0, not an error
