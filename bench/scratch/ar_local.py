"""world=1 run of the fused collective (gather -> reduce -> scatter on local staging only): isolates the
HBM-side cost of the kernel phases from NVLink."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pytorch_distributed_training_tutorials_b200 import group
from pytorch_distributed_training_tutorials_b200.ops import collectives as coll
from pytorch_distributed_training_tutorials_b200.utils.timing import time_cuda
st = group.init(backend="nccl", rank=0, world_size=1)
dev = st.device
arena = coll.make_arena((1 << 30) + (1 << 20), dev)
plain = torch.empty((1 << 30) + (1 << 20), dtype=torch.uint8, device=dev)
for mb in (64, 1024):
    n = (mb << 20) // 4
    x = torch.randn(n, device=dev)
    for blocks in (32, 128, 296):
        table = coll.TensorTable([n], dev, False, False, target_blocks=blocks)
        s = torch.cuda.current_stream(dev); table.bind([x], s)
        for mode in ("oneshot", "twoshot"):
            t = time_cuda(lambda: coll.launch(mode, arena, table, 1.0, s, max_blocks=blocks), iters=5, flush=False)
            passes = 4 if mode == "oneshot" else 6
            print(f"{mb} MiB blocks={blocks} {mode}: {t*1e3:.0f} us  -> {passes * mb / 1024 / (t / 1e3) / 1024:.2f} TiB/s effective over {passes} passes")
    y = torch.empty_like(x)
    t = time_cuda(lambda: y.copy_(x), iters=5, flush=False)
    print(f"{mb} MiB torch copy: {t*1e3:.0f} us -> {2 * mb / 1024 / (t / 1e3) / 1024:.2f} TiB/s")
    sv = arena.buf[: n * 4].view(torch.float32)
    t = time_cuda(lambda: sv.copy_(x), iters=5, flush=False)
    print(f"{mb} MiB torch copy into symmetric buffer: {t*1e3:.0f} us -> {2 * mb / 1024 / (t / 1e3) / 1024:.2f} TiB/s")
group.destroy()
