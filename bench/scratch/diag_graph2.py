import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
from pytorch_distributed_training_tutorials_b200 import group, _lib
from pytorch_distributed_training_tutorials_b200.parallel.ddp import DistributedDataParallel as DDP
import pytorch_distributed_training_tutorials_b200.utils.graph as G
rt = ctypes.CDLL("libcudart.so.12")
def peek(tag):
    print(f"   peek[{tag}] = {rt.cudaPeekAtLastError()}", flush=True)
dev = torch.device("cuda:0")
group.init(backend="nccl", rank=0, world_size=1)
variant = sys.argv[1]
model = resnet50().to(dev); opt = FusedSGD(model.parameters(), lr=1e-3)
if variant != "noddp":
    model = DDP(model, device_ids=[0])
if variant == "nohooks":
    model._reducer.remove_hooks()
    model.rebind_autograd_hooks = lambda: None
x = torch.randn(32, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (32,), device=dev)
peek("setup")
# step-by-step copy of GraphedTrainStep.__init__
cur = torch.cuda.current_stream(dev); side = torch.cuda.Stream(dev); side.wait_stream(cur)
def eager(zero=True):
    if zero: opt.zero_grad(set_to_none=True)
    loss = Fn.cross_entropy(model(x), y); loss.backward(); opt.step(); return loss.detach()
with torch.cuda.stream(side):
    if variant in ("full", "fullitem"): model.rebind_autograd_hooks()
    peek("rebind")
    for i in range(3):
        eager(); peek(f"warmup {i}")
cur.wait_stream(side); torch.cuda.synchronize(); peek("after warmup sync")
g = torch.cuda.CUDAGraph(); opt.zero_grad(set_to_none=True)
with torch.cuda.graph(g):
    l = eager(zero=False)
peek("after capture")
g.replay(); peek("after replay"); torch.cuda.synchronize(); peek("after replay sync")
print("current device", torch.cuda.current_device(), "capturing", torch.cuda.is_current_stream_capturing())
try:
    z = torch.zeros(1, device=dev); print("fresh item", z.item())
except Exception as e:
    print("fresh item FAILED:", str(e).splitlines()[0]); peek("after fail")
try:
    with torch.cuda.stream(side):
        print("item on side stream", (l + 0).item())
except Exception as e:
    print("side item FAILED:", str(e).splitlines()[0])
try:
    print("cpu copy", l.cpu())
except Exception as e:
    print("cpu copy FAILED:", str(e).splitlines()[0])

