import sys, os, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torchvision
import torch.nn.functional as F
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
dev = torch.device("cuda:0")
for (img, bs, lr, ncls) in [(64, 32, 0.01, 10), (64, 32, 0.002, 10), (224, 32, 0.01, 10), (224, 64, 0.01, 1000)]:
    torch.manual_seed(1)
    ref = torchvision.models.resnet50(num_classes=ncls).to(dev).train()
    amp = copy.deepcopy(ref)
    m = resnet50(num_classes=ncls).to(dev).train(); m.load_state_dict(ref.state_dict())
    opts = [FusedSGD(m.parameters(), lr=lr), torch.optim.SGD(ref.parameters(), lr=lr), torch.optim.SGD(amp.parameters(), lr=lr)]
    x = torch.randn(bs, 3, img, img, device=dev); y = torch.randint(0, ncls, (bs,), device=dev)
    L = [[], [], []]
    for it in range(10):
        opts[0].zero_grad(); l = Fn.cross_entropy(m(x), y); l.backward(); opts[0].step(); L[0].append(l.item())
        opts[1].zero_grad(); l = F.cross_entropy(ref(x), y); l.backward(); opts[1].step(); L[1].append(l.item())
        opts[2].zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            l = F.cross_entropy(amp(x), y)
        l.backward(); opts[2].step(); L[2].append(l.item())
    print(f"img{img} bs{bs} lr{lr} C{ncls}")
    for name, l in zip(("mine", "fp32", "amp "), L):
        print("  ", name, " ".join(f"{v:.3f}" for v in l))
