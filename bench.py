#!/usr/bin/env python
"""Headline benchmark: ResNet-50 DDP training throughput (img/s), bf16 compute, synthetic
224x224 ImageNet-shaped batches, random-init weights (BASELINE.json config 3).

    python bench.py --gpus 1 --steps K --warmup W                      # this library
    python bench.py --impl reference --gpus 1 --steps K --warmup W     # unmodified reference stack
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  `value` is the whole-job img/s from CUDA events around exactly K
steps (max over ranks) with inputs resident on the device; `e2e` is the same metric measured
through the public Trainer API including, every step, the pinned-host -> device copy of the batch
and the read-back of the loss.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------- helpers
def dist_env(gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if gpus > 1 and world != gpus:
        raise SystemExit(f"--gpus {gpus} needs one process per GPU: launch with "
                         f"`python -m torch.distributed.run --nproc-per-node {gpus} ... bench.py --gpus {gpus}`")
    return rank, world, local


def max_over_ranks(ms: float, world: int, device) -> float:
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sync_all(world: int, device) -> None:
    if world > 1:
        import torch.distributed as dist
        dist.barrier(device_ids=[device.index])
    torch.cuda.synchronize(device)


def make_host_batches(n: int, batch: int, image: int, classes: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    xs = [torch.randn(batch, 3, image, image, generator=g).pin_memory() for _ in range(n)]
    ys = [torch.randint(0, classes, (batch,), generator=g).pin_memory() for _ in range(n)]
    return xs, ys


def timed_device_steps(run_batch, dev_x, dev_y, steps: int, warmup: int, world: int, device):
    for i in range(warmup):
        run_batch(dev_x[i % len(dev_x)], dev_y[i % len(dev_y)])
    sync_all(world, device)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        run_batch(dev_x[i % len(dev_x)], dev_y[i % len(dev_y)])
    e1.record()
    sync_all(world, device)
    return max_over_ranks(e0.elapsed_time(e1), world, device)


def timed_e2e_steps(step_fn, host_x, host_y, steps: int, warmup: int, world: int, device):
    for i in range(min(warmup, 3)):
        step_fn(host_x[i % len(host_x)], host_y[i % len(host_y)])
    sync_all(world, device)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for i in range(steps):
        last = step_fn(host_x[i % len(host_x)], host_y[i % len(host_y)])
    e1.record()
    sync_all(world, device)
    return max_over_ranks(e0.elapsed_time(e1), world, device), last


def timed_e2e_run(run_batches, host_x, host_y, steps: int, warmup: int, world: int, device):
    """Same contract as ``timed_e2e_steps`` for an API that takes the whole batch stream: every step's
    H2D upload (pinned -> device) and every step's loss read-back happen inside the timed region; the
    library overlaps upload i+1 / read-back i-1 with the kernels of step i."""
    n = len(host_x)
    run_batches([(host_x[i % n], host_y[i % n]) for i in range(min(warmup, 3))])
    sync_all(world, device)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    losses = run_batches((host_x[i % n], host_y[i % n]) for i in range(steps))
    e1.record()
    sync_all(world, device)
    assert len(losses) == steps
    return max_over_ranks(e0.elapsed_time(e1), world, device), losses[-1]


# --------------------------------------------------------------------------- arms
def run_ours(args, rank, world, local):
    from pytorch_distributed_training_tutorials_b200 import _lib, group
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    from pytorch_distributed_training_tutorials_b200.trainer import Trainer

    st = group.init(backend="nccl", rank=rank, world_size=world, local_rank=local)
    device = st.device
    torch.manual_seed(1234)
    model = resnet50(num_classes=args.classes)
    optimizer = FusedSGD(model.parameters(), lr=1e-3)
    trainer = Trainer(model, train_dataloader=None, optimizer=optimizer, gpu_id=local, wrap_ddp=True,
                      ddp_kwargs=dict(allreduce=args.allreduce, comm_dtype=args.comm_dtype,
                                      max_comm_blocks=args.comm_blocks,
                                      gradient_as_bucket_view=not args.no_grad_views,
                                      broadcast_buffers=not args.no_broadcast_buffers),
                      cuda_graph=not args.no_graph)
    host_x, host_y = make_host_batches(2, args.batch_size, args.image, args.classes, seed=rank)
    dev_x = [x.to(device) for x in host_x]
    dev_y = [y.to(device) for y in host_y]

    sampler = ClockSampler(local) if rank == 0 else None
    # warm-up happens inside timed_device_steps; sample clocks only over the timed part
    for i in range(args.warmup):
        trainer._run_batch(dev_x[i % 2], dev_y[i % 2])
    sync_all(world, device)
    if sampler:
        sampler.start()
    def native_launches():
        r = getattr(trainer.model, "_reducer", None)        # collectives the C++ reducer launches itself
        return int(r.stats().get("native_launches", 0)) if r is not None else 0
    launches0 = _lib.launch_count + native_launches()
    ms = timed_device_steps(trainer._run_batch, dev_x, dev_y, args.steps, 0, world, device)
    launches = _lib.launch_count + native_launches() - launches0
    clocks = sampler.stop() if sampler else None
    # both public entry points, every time: the synchronous one shows what the input pipeline costs when
    # nothing hides it, the prefetching one is what a user of the library runs
    sync_ms, sync_loss = timed_e2e_steps(trainer.step, host_x, host_y, args.steps, args.warmup, world, device)
    pre_ms, pre_loss = timed_e2e_run(trainer.run_batches, host_x, host_y, args.steps, args.warmup, world, device)
    e2e_ms, last_loss = (sync_ms, sync_loss) if args.e2e_sync else (pre_ms, pre_loss)
    gb = args.batch_size * world
    info = {"model": "resnet50 (own NHWC bf16 kernels)", "loss": last_loss, "cuda_graph": trainer.cuda_graph,
            "e2e_api": "Trainer.step (synchronous)" if args.e2e_sync else "Trainer.run_batches (prefetching)",
            "e2e_sync": {"api": "Trainer.step (synchronous copy, step, loss.item())",
                         "value": round(gb * args.steps / (sync_ms / 1e3), 1), "ms_per_step": round(sync_ms / args.steps, 3)},
            "e2e_prefetch": {"api": "Trainer.run_batches (upload i+1 / read-back i-1 under step i)",
                             "value": round(gb * args.steps / (pre_ms / 1e3), 1), "ms_per_step": round(pre_ms / args.steps, 3)},
            "ddp": trainer.model.get_logging_data() if hasattr(trainer.model, "get_logging_data") else None}
    group.destroy()
    return ms, e2e_ms, launches, clocks, info


def run_reference(args, rank, world, local):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "ddp_gpus_torchrun.py")):
        src = "/root/reference"
        if os.path.exists(os.path.join(src, "ddp_gpus_torchrun.py")):
            import shutil
            os.makedirs(ref_dir, exist_ok=True)
            for f in ("ddp_gpus.py", "ddp_gpus_torchrun.py"):
                shutil.copy2(os.path.join(src, f), os.path.join(ref_dir, f))
        else:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is empty: the reference is a set of "
                              "scripts without setup.py/pyproject, pip cannot install it and no copy was shipped"}))
            return None
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if "MASTER_PORT" not in os.environ:           # single process without a launcher: any free port
        from pytorch_distributed_training_tutorials_b200.group import find_free_port
        os.environ["MASTER_PORT"] = str(find_free_port()) if world == 1 else "29611"
    os.environ.setdefault("RANK", str(rank)); os.environ.setdefault("WORLD_SIZE", str(world))
    os.environ.setdefault("LOCAL_RANK", str(local))
    sys.path.insert(0, ref_dir)
    import torchvision
    import ddp_gpus_torchrun as ref            # the unmodified reference script (guarded by __main__)

    ref.ddp_setup()                            # init_process_group("nccl") + set_device(LOCAL_RANK)
    device = torch.device("cuda", local)
    torch.manual_seed(1234)

    class Bf16ResNet50(torch.nn.Module):
        """torchvision ResNet-50 run under autocast(bf16): the reference stack's mixed-precision path."""

        def __init__(self):
            super().__init__()
            self.net = torchvision.models.resnet50(num_classes=args.classes)
            self.last_output = None

        def forward(self, x):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = self.net(x)
            self.last_output = out
            return out.float()

    model = Bf16ResNet50()
    optimizer = torch.optim.SGD(model.parameters(), lr=1e-3)
    trainer = ref.Trainer(model=model, train_dataloader=None, optimizer=optimizer)   # .to(gpu) + torch DDP
    host_x, host_y = make_host_batches(2, args.batch_size, args.image, args.classes, seed=rank)
    dev_x = [x.to(device) for x in host_x]
    dev_y = [y.to(device) for y in host_y]

    def e2e_step(xs, ys):
        xs = xs.to(trainer.gpu_id)             # as ddp_gpus_torchrun.py:43-44 (blocking copies)
        ys = ys.to(trainer.gpu_id)
        trainer._run_batch(xs, ys)
        return float(model.last_output.float().mean().item())

    for i in range(args.warmup):
        trainer._run_batch(dev_x[i % 2], dev_y[i % 2])
    sync_all(world, device)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms = timed_device_steps(trainer._run_batch, dev_x, dev_y, args.steps, 0, world, device)
    clocks = sampler.stop() if sampler else None
    e2e_ms, last = timed_e2e_steps(e2e_step, host_x, host_y, args.steps, args.warmup, world, device)
    import torch.distributed as dist
    dist.destroy_process_group()
    return ms, e2e_ms, 0, clocks, {"model": "torchvision resnet50 + autocast(bf16) + torch DDP via reference Trainer",
                                   "result": last}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch-size", type=int, default=256, help="per-GPU batch (weak scaling)")
    ap.add_argument("--image", type=int, default=224)
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--allreduce", default="auto", choices=["auto", "oneshot", "twoshot", "nvls", "nccl"])
    ap.add_argument("--comm-dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of one CUDA graph per step")
    ap.add_argument("--comm-blocks", type=int, default=32, help="CTAs of the fused all-reduce kernel (diagnostics)")
    ap.add_argument("--no-grad-views", action="store_true",
                    help="keep gradients outside the symmetric arena (copy-in / copy-out around the all-reduce)")
    ap.add_argument("--no-broadcast-buffers", action="store_true", help="skip the per-forward buffer broadcast (diagnostics)")
    ap.add_argument("--e2e-sync", action="store_true",
                    help="end-to-end number through the synchronous Trainer.step instead of the prefetching run_batches")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (run it through gpurun)")
    rank, world, local = dist_env(args.gpus)
    res = run_ours(args, rank, world, local) if args.impl == "ours" else run_reference(args, rank, world, local)
    if res is None:
        return
    ms, e2e_ms, launches, clocks, info = res
    if rank != 0:
        return
    global_batch = args.batch_size * world
    value = global_batch * args.steps / (ms / 1e3)
    e2e_value = global_batch * args.steps / (e2e_ms / 1e3)
    in_bytes = args.batch_size * 3 * args.image * args.image * 4 + args.batch_size * 8
    out = {
        "metric": "ResNet-50 DDP training throughput", "value": round(value, 1), "unit": "img/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic (random-init weights, randn images, random labels)",
        "impl": args.impl,
        "config": {"model": "resnet50", "global_batch": global_batch, "per_gpu_batch": args.batch_size,
                   "image": f"3x{args.image}x{args.image}", "seq_len": None, "parallelism": f"dp{world}",
                   "optimizer": "SGD lr=1e-3", "loss": "cross_entropy",
                   "l2_policy": "inputs larger than L2 (154 MB fp32 batch per step, 2 alternating batches; "
                                "activations >> 126 MB)",
                   "allreduce": args.allreduce, "comm_dtype": args.comm_dtype},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "img/s", "ms_per_step": round(e2e_ms / args.steps, 3),
                "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": 4},
        "gpu_launches": launches,
        "baseline_note": "the reference publishes no number for this config (BASELINE.md section 2); "
                         "compare with `--impl reference` on the same box",
        "info": info,
    }
    print(json.dumps(out))


if __name__ == "__main__":
    main()
