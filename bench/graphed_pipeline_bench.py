"""Single-process 2-GPU ResNet-50: naive split vs eager pipeline vs CUDA-graph replayed pipeline.

    python bench/graphed_pipeline_bench.py [--batch 512 --chunks 1,2,4,8 --cut layer1,layer2,balanced,layer2.1]

A cut is any unit of ``ResNet.unit_names()`` (whole layers or single bottleneck blocks); ``balanced`` lets
``ModelParallelResNet50.balanced`` measure the units and pick the cut.  ``--no-eager`` skips the eager variants.

Every variant runs zero_grad + forward + loss + backward + FusedSGD step on the same device-resident
mini-batch; device-timed with CUDA events on cuda:0 after synchronising both GPUs. BASELINE.json config 4.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.utils.graph import GraphedTrainStep


def sync_all():
    for d in range(torch.cuda.device_count()):
        torch.cuda.synchronize(d)


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    sync_all()
    with torch.cuda.device(0):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
    for _ in range(steps):
        fn()
    sync_all()
    with torch.cuda.device(0):
        e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--size", type=int, default=224)
    ap.add_argument("--chunks", type=str, default="1,2,4,8")
    ap.add_argument("--cut", type=str, default="layer1,layer2")
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--no-eager", action="store_true")
    args = ap.parse_args()
    B = args.batch
    two = torch.cuda.device_count() >= 2
    d0, d1 = "cuda:0", ("cuda:1" if two else "cuda:0")
    x = torch.randn(B, 3, args.size, args.size, device=d0)
    y = torch.randint(0, 1000, (B,), device=d1)
    res = {"batch": B, "size": args.size, "devices": [d0, d1]}

    m = resnet50().to(d0); o = b2.FusedSGD(m.parameters(), lr=1e-3); y0 = y.to(d0)
    def single():
        o.zero_grad(); Fn.cross_entropy(m(x), y0).backward(); o.step()
    res["single_gpu_eager_ms"] = timed(single, args.steps)
    g = GraphedTrainStep(m, Fn.cross_entropy, o, x, y0)
    res["single_gpu_graph_ms"] = timed(lambda: g(x, y0), args.steps)
    del m, o, g
    torch.cuda.empty_cache()
    print(json.dumps(res), flush=True)

    for cut in args.cut.split(","):
        if cut != "balanced" and not args.no_eager:
            mp = ModelParallelResNet50(dev0=d0, dev1=d1, split_after=cut); o = b2.FusedSGD(mp.parameters(), lr=1e-3)
            def naive():
                o.zero_grad(); Fn.cross_entropy(mp(x), y).backward(); o.step()
            res[f"cut_{cut}/naive_eager_ms"] = timed(naive, args.steps)
            del mp, o
        for ch in [int(c) for c in args.chunks.split(",")]:
            if B % ch:
                continue
            if cut == "balanced":
                pm = PipelineParallelResNet50.balanced(x, chunks=ch, split_size=B // ch, dev0=d0, dev1=d1)
                res[f"cut_balanced/chunks{ch}_chosen"] = pm.split_after
                res[f"cut_balanced/chunks{ch}_unit_ms"] = {k: round(v, 3) for k, v in pm.unit_times_ms.items()}
                print(f"balanced cut for {ch} chunks: after {pm.split_after}", flush=True)
            else:
                pm = PipelineParallelResNet50(split_size=B // ch, dev0=d0, dev1=d1, split_after=cut)
            o = b2.FusedSGD(pm.parameters(), lr=1e-3)
            if ch > 1 and not args.no_eager:
                def eager():
                    o.zero_grad(); Fn.cross_entropy(pm(x), y).backward(); o.step()
                res[f"cut_{cut}/pipeline_eager_chunks{ch}_ms"] = timed(eager, args.steps)
            pipe = pm.graphed(Fn.cross_entropy, x, y)
            def graphed():
                pipe.zero_grad(); pipe.train_step(x, y); o.step()
            res[f"cut_{cut}/pipeline_graphed_chunks{ch}_ms"] = timed(graphed, args.steps)
            print(f"cut {cut} chunks {ch}: graphed {res[f'cut_{cut}/pipeline_graphed_chunks{ch}_ms']:.1f} ms", flush=True)
            del pipe, pm, o
            torch.cuda.empty_cache()
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/graphed_pipeline_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
