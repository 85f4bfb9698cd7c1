"""cProfile of the notebook's train() workload on this library (eager, single GPU): where the host time goes."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
sys.path.insert(0, os.path.join(sys.path[0], 'bench'))
from mp_bench import train_ours, num_classes
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD

m = resnet50(num_classes=num_classes).to("cuda:0"); o = FusedSGD(m.parameters(), lr=1e-3)
for _ in range(3):
    train_ours(m, "cuda:0", "cuda:0", o)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    train_ours(m, "cuda:0", "cuda:0", o)
torch.cuda.synchronize()
print(f"train(): {(time.perf_counter() - t0) / 5 * 1e3:.1f} ms wall")
t0 = time.perf_counter()
for _ in range(15):
    torch.randn(120, 3, 128, 128)
print(f"host randn of one batch: {(time.perf_counter() - t0) / 15 * 1e3:.2f} ms")
pr = cProfile.Profile(); pr.enable()
for _ in range(5):
    train_ours(m, "cuda:0", "cuda:0", o)
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22); print(s.getvalue()[:5000])
