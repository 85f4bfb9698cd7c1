"""Toy MLP throughput: DataParallel (single process) vs DDP (one process per GPU), this library vs torch.
BASELINE.json config 2.  Synthetic regression data, MSE loss, SGD; device-timed, max over ranks.

    python bench/toy_bench.py --mode dp --gpus 8
    python -m torch.distributed.run --nproc-per-node 8 bench/toy_bench.py --mode ddp
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn
import torch.nn.functional as F

IN, HID, DEPTH, OUT = 1024, 4096, 4, 8


def torch_mlp():
    dims = [IN] + [HID] * DEPTH
    layers = []
    for a, b in zip(dims[:-1], dims[1:]):
        layers += [nn.Linear(a, b), nn.ReLU()]
    layers.append(nn.Linear(HID, OUT))
    return nn.Sequential(*layers)


class Autocast(nn.Module):
    def __init__(self, m):
        super().__init__(); self.m = m

    def forward(self, x):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return self.m(x).float()


def timed(step, steps, warmup, dev, dist_on):
    for _ in range(warmup):
        step()
    if dist_on:
        import torch.distributed as dist
        dist.barrier(device_ids=[dev.index])
    torch.cuda.synchronize(dev)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record(); torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    if dist_on:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev, dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
    return ms / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["dp", "ddp"], required=True)
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--per-gpu-batch", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=8)
    args = ap.parse_args()
    import pytorch_distributed_training_tutorials_b200 as b2
    from pytorch_distributed_training_tutorials_b200.models.toy import MLP
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    res = {"mode": args.mode, "model": f"MLP {IN}->{HID}x{DEPTH}->{OUT}", "per_gpu_batch": args.per_gpu_batch}
    if args.mode == "dp":
        n = args.gpus or torch.cuda.device_count()
        dev = torch.device("cuda:0")
        B = args.per_gpu_batch * n
        x = torch.randn(B, IN, device=dev); y = torch.randn(B, OUT, device=dev)
        ids = list(range(n))
        mine = MLP(IN, HID, DEPTH, OUT).to(dev)
        dp = b2.DataParallel(mine, device_ids=ids) if n > 1 else mine
        opt = b2.FusedSGD(mine.parameters(), lr=1e-4)

        def step_mine():
            opt.zero_grad(); Fn.mse_loss(dp(x), y).backward(); opt.step()
        ms = timed(step_mine, args.steps, args.warmup, dev, False)
        res["ours_ms"] = ms; res["ours_samples_per_s"] = B / ms * 1e3
        if n > 1:
            # persistent replicas + one CUDA graph per device (parallel/graphed_data_parallel.py)
            gm = MLP(IN, HID, DEPTH, OUT).to(dev)
            gopt = b2.FusedSGD(gm.parameters(), lr=1e-4)
            gdp = b2.GraphedDataParallel(gm, ids, Fn.mse_loss, x, y)

            def step_graphed():
                gopt.zero_grad(); gdp.train_step(x, y); gopt.step()
            ms = timed(step_graphed, args.steps, args.warmup, dev, False)
            res["ours_graphed_ms"] = ms; res["ours_graphed_samples_per_s"] = B / ms * 1e3
            del gdp, gm, gopt
        ref_m = Autocast(torch_mlp()).to(dev)
        rdp = nn.DataParallel(ref_m, device_ids=ids) if n > 1 else ref_m
        ropt = torch.optim.SGD(ref_m.parameters(), lr=1e-4)

        def step_ref():
            ropt.zero_grad(); F.mse_loss(rdp(x), y).backward(); ropt.step()
        ms = timed(step_ref, args.steps, args.warmup, dev, False)
        res["torch_ms"] = ms; res["torch_samples_per_s"] = B / ms * 1e3
        res["n_gpus"] = n
        rank = 0
    else:
        st = b2.init(backend="nccl")
        dev, n, rank = st.device, st.world_size, st.rank
        B = args.per_gpu_batch
        x = torch.randn(B, IN, device=dev); y = torch.randn(B, OUT, device=dev)
        mine = b2.DistributedDataParallel(MLP(IN, HID, DEPTH, OUT).to(dev), device_ids=[dev.index],
                                          gradient_as_bucket_view=True)
        opt = b2.FusedSGD(mine.parameters(), lr=1e-4)
        from pytorch_distributed_training_tutorials_b200.utils.graph import GraphedTrainStep
        g = GraphedTrainStep(mine, Fn.mse_loss, opt, x, y)
        ms = timed(lambda: g(x, y), args.steps, args.warmup, dev, n > 1)
        res["ours_ms"] = ms; res["ours_samples_per_s"] = B * n / ms * 1e3
        mine2 = b2.DistributedDataParallel(MLP(IN, HID, DEPTH, OUT).to(dev), device_ids=[dev.index])
        opt2 = b2.FusedSGD(mine2.parameters(), lr=1e-4)

        def step_eager():
            opt2.zero_grad(); Fn.mse_loss(mine2(x), y).backward(); opt2.step()
        ms = timed(step_eager, args.steps, args.warmup, dev, n > 1)
        res["ours_eager_ms"] = ms; res["ours_eager_samples_per_s"] = B * n / ms * 1e3
        ref = nn.parallel.DistributedDataParallel(Autocast(torch_mlp()).to(dev), device_ids=[dev.index])
        ropt = torch.optim.SGD(ref.parameters(), lr=1e-4)

        def step_ref():
            ropt.zero_grad(); F.mse_loss(ref(x), y).backward(); ropt.step()
        ms = timed(step_ref, args.steps, args.warmup, dev, n > 1)
        res["torch_ms"] = ms; res["torch_samples_per_s"] = B * n / ms * 1e3
        res["n_gpus"] = n
    if rank == 0:
        print(json.dumps(res))
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(res, open(f"gpurun_out/toy_bench_{args.mode}_{res['n_gpus']}.json", "w"), indent=1)
    if args.mode == "ddp":
        b2.destroy()


if __name__ == "__main__":
    main()
