"""Multi-process pipelined model parallelism (one process per stage, NCCL p2p) on ResNet-50.

    python -m torch.distributed.run --nproc-per-node 2 bench/pipeline_bench.py [--batch 512 --chunks 4 --cut layer1]

Rank 0 first times the same mini-batch on ONE GPU (whole model), then both ranks time the 2-stage
pipeline; device-timed, max over ranks. Naive (one micro-batch) = --chunks 1.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch import nn
import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50, ResNet
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.parallel.pipeline import PipelineStage


class Stage(nn.Module):
    def __init__(self, net: ResNet, first: str, last: str, head: bool):
        super().__init__()
        self.net, self.first, self.last, self.head = net, first, last, head

    def forward(self, x):
        x = self.net.run_segments(x, self.first, self.last)
        return self.net.fc(x) if self.head else x


def timed(fn, steps, warmup, dev, world):
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier(device_ids=[dev.index])
    torch.cuda.synchronize(dev)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--chunks", type=str, default="1,2,4,8")
    ap.add_argument("--cut", type=str, default="layer1,layer2")
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    st = b2.init(backend="nccl")
    dev, rank, world = st.device, st.rank, st.world_size
    assert world == 2
    B = args.batch
    x = torch.randn(B, 3, 224, 224, device=dev) if rank == 0 else None
    y = torch.randint(0, 1000, (B,), device=dev)
    res = {"batch": B}
    if rank == 0:
        m = resnet50().to(dev); o = b2.FusedSGD(m.parameters(), lr=1e-3)
        def single():
            o.zero_grad(); Fn.cross_entropy(m(x), y).backward(); o.step()
        res["single_gpu_ms"] = timed(single, args.steps, 3, dev, 1)
        del m, o
        torch.cuda.empty_cache()
    dist.barrier(device_ids=[dev.index])
    for cut in args.cut.split(","):
        nxt = ResNet.SEGMENTS[ResNet.SEGMENTS.index(cut) + 1]
        torch.manual_seed(0)
        net = resnet50()
        mod = Stage(net, "stem", cut, False) if rank == 0 else Stage(net, nxt, "pool", True)
        # only this stage's parameters go to the GPU / optimizer
        mine = [p for n, p in net.named_parameters()]
        net.to(dev)
        opt = b2.FusedSGD(net.parameters(), lr=1e-3)
        for ch in [int(c) for c in args.chunks.split(",")]:
            stage = PipelineStage(mod, chunks=ch)
            def step():
                opt.zero_grad()
                stage.train_step(x=x, target=y if rank == 1 else None, loss_fn=Fn.cross_entropy if rank == 1 else None, n_rows=B)
                opt.step()
            ms = timed(step, args.steps, 3, dev, world)
            res[f"pipeline_cut_{cut}_chunks{ch}_ms"] = ms
            if rank == 0:
                print(f"cut after {cut}, {ch} micro-batch(es): {ms:.1f} ms/step  ({B / ms * 1e3:.0f} img/s)", flush=True)
        del net, opt
        torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(res))
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(res, open("gpurun_out/pipeline_bench.json", "w"), indent=1)
    b2.destroy()


if __name__ == "__main__":
    main()
