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
        for m in (["twoshot", "nvls"] if arena.has_multicast else ["twoshot"]):
            row[m + "_inplace_ms"] = timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=args.blocks,
                                                               in_place=True), iters, dev)
        if size >= (1 << 20):
            table = coll.TensorTable([n], dev, True, False, target_blocks=args.blocks)
            table.bind([x], stream)
            for m in (["twoshot", "nvls"] if arena.has_multicast else ["twoshot"]):
                row[m + "_bf16wire_ms"] = timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=args.blocks), iters, dev)
        arena.check_status()
        rows.append(row)
        size *= 4
    probe = {}
    if arena.has_multicast and max_bytes >= (64 << 20):
        # how many CTAs does the NVLS kernel want?  (64 MiB fp32 in place)
        n = (64 << 20) // 4
        stream = torch.cuda.current_stream(dev)
        for blocks in (16, 32, 64, 128, 256):
            table = coll.TensorTable([n], dev, False, False, target_blocks=blocks)
            (xv,) = coll.stage_views(arena, table, [(n,)])
            table.bind([xv], stream)
            probe[blocks] = {m: timed(lambda: coll.launch(m, arena, table, 1.0, stream, max_blocks=blocks, in_place=True), 10, dev)
                             for m in ("twoshot", "nvls")}
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
        # roofline: bytes every GPU must push through its 900 GB/s/direction NVLink port
        #   two-shot: 2 S (N-1)/N   (== busbw definition)      NVLS: S (N+1)/N   (switch reduces / fans out)
        R = 900e9
        print()
        print("| bytes | best in-place kernel | us | % of link bound (900 GB/s) | % of measured peer-copy rate (770 GB/s) | vs NCCL |")
        print("|---|---|---|---|---|---|")
        for r in rows:
            cands = {k: r[k] for k in ("twoshot_inplace_ms", "nvls_inplace_ms") if k in r}
            k = min(cands, key=cands.get)
            t = cands[k] * 1e-3
            need = r["bytes"] * ((world + 1) / world if k.startswith("nvls") else f)
            print(f"| {r['bytes']} | {k[:-11]} | {t * 1e6:.1f} | {100 * need / R / t:.1f} | {100 * need / 770e9 / t:.1f} | "
                  f"{r['nccl_ms'] * 1e-3 / t:.2f}x |")
        if probe:
            print()
            print("64 MiB fp32 in place, ms by CTA count: " + "; ".join(
                f"{b}: twoshot {v['twoshot']:.3f} nvls {v['nvls']:.3f}" for b, v in probe.items()))
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump({"world": world, "cta_probe_64MiB_ms": probe, "multicast": arena.has_multicast, "arena": arena.flavour,
                   "multicast_error": arena.multicast_error, "rows": rows},
                  open(f"gpurun_out/allreduce_sweep_{world}.json", "w"), indent=1)
    group.destroy()


if __name__ == "__main__":
    main()
