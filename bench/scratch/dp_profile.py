"""Host-side profile + device timeline split of one DataParallel MLP step (diagnostics)."""
import sys, os, cProfile, pstats, io, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import pytorch_distributed_training_tutorials_b200 as b2
from pytorch_distributed_training_tutorials_b200.models.toy import MLP
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.parallel import data_parallel as DPm

n = torch.cuda.device_count()
dev = torch.device("cuda:0")
B = 2048 * n
x = torch.randn(B, 1024, device=dev); y = torch.randn(B, 8, device=dev)
m = MLP(1024, 4096, 4, 8).to(dev)
dp = b2.DataParallel(m, device_ids=list(range(n)))
opt = b2.FusedSGD(m.parameters(), lr=1e-4)
def step():
    opt.zero_grad(); Fn.mse_loss(dp(x), y).backward(); opt.step()
for _ in range(8): step()
for d in range(n): torch.cuda.synchronize(d)
# phase timing (host wall clock with syncs): forward pieces
def sync():
    for d in range(n): torch.cuda.synchronize(d)
for name, fn in [("scatter", lambda: DPm.scatter_kwargs((x,), None, list(range(n)))),
                 ("replicate", lambda: DPm.replicate(m, list(range(n))))]:
    sync(); t0 = time.perf_counter()
    for _ in range(10): out = fn()
    t1 = time.perf_counter(); sync(); t2 = time.perf_counter()
    print(f"{name}: host {1e3*(t1-t0)/10:.3f} ms/iter, +drain {1e3*(t2-t1)/10:.3f}")
sync(); t0 = time.perf_counter()
for _ in range(10): step()
t1 = time.perf_counter(); sync(); t2 = time.perf_counter()
print(f"step: host {1e3*(t1-t0)/10:.3f} ms/iter, total {1e3*(t2-t0)/10:.3f}")
pr = cProfile.Profile(); pr.enable()
for _ in range(10): step()
sync(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(30); print(s.getvalue()[:5500])
