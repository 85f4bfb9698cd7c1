"""Where does a graph-replayed pipeline step go?  Times every (stage, slot) graph of a GraphedPipeline on its own
(forward and backward), the boundary copies, zero_grad and the optimizer step, then prints the fill-drain critical
path those numbers predict next to the measured step.

    python bench/diag_pipeline_stages.py [--batch 512 --chunks 4 --cut layer2.1]
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.models.resnet_mp import PipelineParallelResNet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn


def sync_all():
    for d in range(torch.cuda.device_count()):
        torch.cuda.synchronize(d)


def time_on(dev, stream, fn, reps=5):
    with torch.cuda.device(dev):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(reps + 1):
            sync_all()
            with torch.cuda.stream(stream):
                e0.record(stream); fn(); e1.record(stream)
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--cut", type=str, default="layer2.1")
    args = ap.parse_args()
    B, m = args.batch, args.chunks
    two = torch.cuda.device_count() >= 2
    d0, d1 = "cuda:0", ("cuda:1" if two else "cuda:0")
    x = torch.randn(B, 3, 224, 224, device=d0); y = torch.randint(0, 1000, (B,), device=d1)
    pm = PipelineParallelResNet50(split_size=B // m, dev0=d0, dev1=d1, split_after=args.cut)
    o = b2.FusedSGD(pm.parameters(), lr=1e-3)
    pipe = pm.graphed(Fn.cross_entropy, x, y)
    def step():
        pipe.zero_grad(); pipe.train_step(x, y); o.step()
    for _ in range(3):
        step()
    sync_all()
    res = {"cut": args.cut, "chunks": m}
    with torch.cuda.device(0):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(8):
            step()
        sync_all(); e1.record(); e1.synchronize()
    res["step_ms"] = e0.elapsed_time(e1) / 8
    devs, st = pipe.devices, pipe.streams
    f = [[time_on(devs[k], st[k], pipe.fwd[k][i].replay) for i in range(m)] for k in range(2)]
    b = [[time_on(devs[k], st[k], pipe.bwd[k][i].replay) for i in range(m)] for k in range(2)]
    res["fwd_ms"], res["bwd_ms"] = f, b
    res["copy_act_ms"] = time_on(devs[1], st[1], lambda: pipe.sin[1][0].detach().copy_(pipe.out[0][0], non_blocking=True))
    res["copy_grad_ms"] = time_on(devs[0], st[0], lambda: pipe.gout[0][0].copy_(pipe.gin[1][0], non_blocking=True))
    res["zero_grad_ms"] = time_on(devs[0], torch.cuda.current_stream(devs[0]), pipe.zero_grad)
    res["opt_step_ms"] = time_on(devs[0], torch.cuda.current_stream(devs[0]), o.step)
    res["input_copy_ms"] = time_on(devs[0], st[0], lambda: [pipe.sin[0][i].copy_(x[i * (B // m):(i + 1) * (B // m)]) for i in range(m)])
    # fill-drain critical path from the per-graph times
    t0 = t1 = 0.0
    endf0 = []
    for i in range(m):
        t0 += f[0][i]; endf0.append(t0)
    for i in range(m):
        t1 = max(t1, endf0[i] + res["copy_act_ms"]) + f[1][i]
    endb1 = []
    for i in reversed(range(m)):
        t1 += b[1][i]; endb1.append(t1)
    for j, i in enumerate(reversed(range(m))):
        t0 = max(t0, endb1[j] + res["copy_grad_ms"]) + b[0][i]
    res["predicted_ms"] = max(t0, t1) + res["zero_grad_ms"] + res["opt_step_ms"] + res["input_copy_ms"]
    res["stage_busy_ms"] = [sum(f[0]) + sum(b[0]), sum(f[1]) + sum(b[1])]
    print(json.dumps(res, indent=1))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/pipe_stages_{args.cut}_{m}.json", "w"), indent=1)


if __name__ == "__main__":
    main()
