"""All-reduce bandwidth sweep 1 KB - 1 GB: this library's one-shot / two-shot / NVLS kernels vs NCCL.

    python -m torch.distributed.run --nproc-per-node N bench/allreduce_sweep.py [--max-mb 1024]

Device-timed with CUDA events, max over ranks; prints a markdown table on rank 0 and writes
gpurun_out/allreduce_sweep_N.json (BASELINE.json config 5).  algbw = S / t, busbw = algbw * 2(N-1)/N.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from pytorch_distributed_training_tutorials_b200 import group
from pytorch_distributed_training_tutorials_b200.ops import collectives as coll


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(dev); dist.barrier(device_ids=[dev.index])
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--blocks", type=int, default=128)
    args = ap.parse_args()
    st = group.init(backend="nccl")
    dev, world, rank = st.device, st.world_size, st.rank
    max_bytes = args.max_mb << 20
    arena = coll.make_arena(max_bytes + (1 << 20), dev)
    modes = ["oneshot", "twoshot"] + (["nvls"] if arena.has_multicast else [])
    rows = []
    size = 1024
    while size <= max_bytes:
        n = size // 4
        x = torch.randn(n, device=dev)
        iters = 200 if size <= (1 << 20) else 30 if size <= (64 << 20) else 8
        row = {"bytes": size}
        row["nccl_ms"] = timed(lambda: dist.all_reduce(x), iters, dev)
        stream = torch.cuda.current_stream(dev)
        for m in modes:
            if m == "oneshot" and size > (8 << 20):
                continue
            table = coll.TensorTable([n], dev, False, False, target_blocks=args.blocks)
            table.bind([x], stream)
            row[m + "_ms"] = timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=args.blocks), iters, dev)
        # tensors that live in the symmetric arena (the DDP gradient-view mode): no copy-in / copy-out
        table = coll.TensorTable([n], dev, False, False, target_blocks=args.blocks)
        (xv,) = coll.stage_views(arena, table, [(n,)])
        xv.copy_(x)
        table.bind([xv], stream)
        m = "nvls" if arena.has_multicast else "twoshot"
        row[m + "_inplace_ms"] = timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=args.blocks,
                                                           in_place=True), iters, dev)
        if size >= (1 << 20):
            table = coll.TensorTable([n], dev, True, False, target_blocks=args.blocks)
            table.bind([x], stream)
            m = "nvls" if arena.has_multicast else "twoshot"
            row[m + "_bf16wire_ms"] = timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=args.blocks), iters, dev)
        arena.check_status()
        rows.append(row)
        size *= 4
    if rank == 0:
        f = 2 * (world - 1) / world
        keys = [k for k in ("nccl_ms", "oneshot_ms", "twoshot_ms", "nvls_ms", "twoshot_inplace_ms", "nvls_inplace_ms",
                            "twoshot_bf16wire_ms", "nvls_bf16wire_ms") if any(k in r for r in rows)]
        print("| bytes | " + " | ".join(k[:-3] + " us / busbw GB/s" for k in keys) + " |")
        print("|---|" + "---|" * len(keys))
        for r in rows:
            cells = []
            for k in keys:
                if k in r:
                    cells.append(f"{r[k] * 1e3:.1f} / {r['bytes'] / (r[k] * 1e-3) * f / 1e9:.1f}")
                else:
                    cells.append("-")
            print(f"| {r['bytes']} | " + " | ".join(cells) + " |")
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump({"world": world, "multicast": arena.has_multicast, "rows": rows}, open(f"gpurun_out/allreduce_sweep_{world}.json", "w"), indent=1)
    group.destroy()


if __name__ == "__main__":
    main()
