"""Host-side cost of one training step: tiny batch so the GPU is never the bottleneck."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
dev = torch.device("cuda:0")
m = resnet50().to(dev); opt = FusedSGD(m.parameters(), lr=1e-3)
x = torch.randn(8, 3, 64, 64, device=dev); y = torch.randint(0, 1000, (8,), device=dev)
def step():
    opt.zero_grad(); loss = Fn.cross_entropy(m(x), y); loss.backward(); opt.step()
for _ in range(5): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): step()
torch.cuda.synchronize()
print(f"host-bound step: {(time.perf_counter() - t0) / 20 * 1e3:.2f} ms")
t0 = time.perf_counter()
for _ in range(20):
    out = m(x)
torch.cuda.synchronize()
print(f"host-bound forward only: {(time.perf_counter() - t0) / 20 * 1e3:.2f} ms")
import torchvision
r = torchvision.models.resnet50().to(dev); ro = torch.optim.SGD(r.parameters(), lr=1e-3)
def rstep():
    ro.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        l = torch.nn.functional.cross_entropy(r(x), y)
    l.backward(); ro.step()
for _ in range(5): rstep()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): rstep()
torch.cuda.synchronize()
print(f"torchvision autocast host-bound step: {(time.perf_counter() - t0) / 20 * 1e3:.2f} ms")
