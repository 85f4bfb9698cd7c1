import sys, os, cProfile, pstats, io
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
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28); print(s.getvalue()[:6000])
