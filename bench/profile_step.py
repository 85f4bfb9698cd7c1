"""One profiled ResNet-50 training step (batch 256, own kernels) for ncu: warm up, then bracket exactly
one step with cudaProfilerStart/Stop (run ncu with --profile-from-start off)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD

bs = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = resnet50().to(dev)
opt = FusedSGD(m.parameters(), lr=1e-3)
x = torch.randn(bs, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (bs,), device=dev)


def step():
    opt.zero_grad()
    loss = Fn.cross_entropy(m(x), y)
    loss.backward()
    opt.step()
    return loss


for _ in range(4):
    step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    step()
e1.record(); torch.cuda.synchronize()
print(f"bs={bs} ms/step={e0.elapsed_time(e1) / 5:.2f}")
