"""Does the two-stage pipeline overlap? Per-device busy time and cross-device overlap of one step,
from torch.profiler kernel timestamps (diagnostic only: numbers under a profiler are not bench values)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import profile, ProfilerActivity
from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
B, split = 1024, 256
x = torch.randn(B, 3, 224, 224, device="cuda:0"); y = torch.randn(B, 1000, device="cuda:1")
for name, m in (("naive", ModelParallelResNet50()), ("pipelined", PipelineParallelResNet50(split_size=split))):
    o = FusedSGD(m.parameters(), lr=1e-3)
    def step():
        o.zero_grad(); Fn.mse_loss(m(x), y).backward(); o.step()
    for _ in range(3): step()
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step(); torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    iv = {0: [], 1: []}
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA and e.device_index in iv:
            iv[e.device_index].append((e.time_range.start, e.time_range.end))
    def union(v):
        v = sorted(v); out = []
        for s, t in v:
            if out and s <= out[-1][1]: out[-1][1] = max(out[-1][1], t)
            else: out.append([s, t])
        return out
    u0, u1 = union(iv[0]), union(iv[1])
    busy0 = sum(t - s for s, t in u0); busy1 = sum(t - s for s, t in u1)
    # overlap
    i = j = 0; ov = 0
    while i < len(u0) and j < len(u1):
        s = max(u0[i][0], u1[j][0]); t = min(u0[i][1], u1[j][1])
        if t > s: ov += t - s
        if u0[i][1] < u1[j][1]: i += 1
        else: j += 1
    start = min(u0[0][0], u1[0][0]); end = max(u0[-1][1], u1[-1][1])
    print(f"{name}: span {(end - start) / 1e3:.1f} ms | gpu0 busy {busy0 / 1e3:.1f} ms | gpu1 busy {busy1 / 1e3:.1f} ms | both busy {ov / 1e3:.1f} ms")
    del m, o
