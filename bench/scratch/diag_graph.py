import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops import nn as bnn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
from pytorch_distributed_training_tutorials_b200.utils.graph import GraphedTrainStep
dev = torch.device("cuda:0")

def check(tag):
    try:
        torch.cuda.synchronize()
        t = torch.zeros(1, device=dev); t.item()
        print("ok  ", tag, flush=True)
    except Exception as e:
        print("FAIL", tag, str(e).splitlines()[0], flush=True)
        raise

which = sys.argv[1] if len(sys.argv) > 1 else "mlp"
torch.manual_seed(0)
if which == "mlp":
    model = torch.nn.Sequential(bnn.Linear(64, 128, relu=True, out_fp32=False), bnn.Linear(128, 8)).to(dev)
    x = torch.randn(256, 64, device=dev); y = torch.randn(256, 8, device=dev); loss_fn = Fn.mse_loss
elif which == "convbn":
    class M(torch.nn.Module):
        def __init__(s):
            super().__init__(); s.c = bnn.Conv2d(64, 64, 3); s.b = bnn.BatchNorm2d(64, relu=True); s.p = bnn.GlobalAvgPool(); s.f = bnn.Linear(64, 8)
        def forward(s, x): return s.f(s.p(s.b(s.c(x))))
    model = M().to(dev); x = torch.randn(16, 56, 56, 64, device=dev).bfloat16(); y = torch.randn(16, 8, device=dev); loss_fn = Fn.mse_loss
else:
    model = resnet50().to(dev); x = torch.randn(32, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (32,), device=dev); loss_fn = Fn.cross_entropy
opt = FusedSGD(model.parameters(), lr=1e-3)
check("setup")
g = GraphedTrainStep(model, loss_fn, opt, x, y)
check("captured")
for i in range(3):
    l = g(x, y)
    check(f"replay {i}")
    print("   loss", float(l))

if which == "resnet" and len(sys.argv) > 2:
    # the bench flow: Trainer + DDP(world=1) + alternating device batches + host batches
    from pytorch_distributed_training_tutorials_b200 import group
    from pytorch_distributed_training_tutorials_b200.trainer import Trainer
    group.init(backend="nccl", rank=0, world_size=1)
    model2 = resnet50().to(dev); opt2 = FusedSGD(model2.parameters(), lr=1e-3)
    tr = Trainer(model2, None, opt2, gpu_id=0, wrap_ddp=True, cuda_graph=True)
    xs = [torch.randn(32, 3, 224, 224, device=dev) for _ in range(2)]; ys = [torch.randint(0, 1000, (32,), device=dev) for _ in range(2)]
    check("trainer setup")
    for i in range(4):
        tr._run_batch(xs[i % 2], ys[i % 2]); check(f"run_batch {i}")
    hx = torch.randn(32, 3, 224, 224).pin_memory(); hy = torch.randint(0, 1000, (32,)).pin_memory()
    for i in range(3):
        print("   e2e loss", tr.step(hx, hy)); check(f"step {i}")
    group.destroy()
