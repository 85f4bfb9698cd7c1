"""Multi-process pipeline parallelism on CPU (gloo, 2 and 3 stages): gradients equal the unsplit model's."""
import os

import torch
import torch.distributed as dist

from pytorch_distributed_training_tutorials_b200 import group
from pytorch_distributed_training_tutorials_b200.launch.spawn import spawn
from pytorch_distributed_training_tutorials_b200.parallel.pipeline import PipelineStage
from pytorch_distributed_training_tutorials_b200.parallel.model_parallel import split_sequential


def _full():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16), torch.nn.Tanh(),
                               torch.nn.Linear(16, 8), torch.nn.Tanh(), torch.nn.Linear(8, 3))


def _worker(rank, world, port, out_dir):
    group.init(backend="gloo", rank=rank, world_size=world, master_addr="127.0.0.1", master_port=port)
    full = _full()
    cuts = {2: [4], 3: [2, 4]}[world]
    stage_mod = split_sequential(full, cuts)[rank]
    stage = PipelineStage(stage_mod, chunks=4)
    g = torch.Generator().manual_seed(3)
    ref = _full()
    for step in range(2):
        x = torch.randn(10, 6, generator=g); y = torch.randn(10, 3, generator=g)
        for p in stage_mod.parameters():
            p.grad = None
        loss = stage.train_step(x=x if rank == 0 else None, target=y if rank == world - 1 else None,
                                loss_fn=torch.nn.functional.mse_loss if rank == world - 1 else None, n_rows=10)
        ref.zero_grad()
        rl = torch.nn.functional.mse_loss(ref(x), y); rl.backward()
        ref_stage = split_sequential(ref, cuts)[rank]
        for p, q in zip(stage_mod.parameters(), ref_stage.parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-6), (rank, step, (p.grad - q.grad).abs().max())
        if rank == world - 1:
            assert abs(loss.item() - rl.item()) < 1e-6
    out = stage(x if rank == 0 else None)
    if rank == world - 1:
        assert torch.allclose(out, full(x), atol=1e-6)
    group.destroy()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def test_pipeline_two_stages(tmp_path):
    spawn(_worker, args=(2, group.find_free_port(), str(tmp_path)), nprocs=2)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_pipeline_three_stages(tmp_path):
    spawn(_worker, args=(3, group.find_free_port(), str(tmp_path)), nprocs=3)
    assert all((tmp_path / f"ok{r}").exists() for r in range(3))
