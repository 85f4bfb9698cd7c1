"""Gradient agreement of the own-kernel ResNet-50 with torchvision fp32, with torchvision-under-autocast
as the yardstick for what bf16 costs on the same batch."""
import sys, os, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torchvision
import torch.nn.functional as F
from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn

dev = torch.device("cuda:0")
for bs in (16, 64):
    torch.manual_seed(0)
    ref = torchvision.models.resnet50(num_classes=1000).to(dev).train()
    amp = copy.deepcopy(ref)
    mine = resnet50(num_classes=1000).to(dev).train()
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(bs, 3, 224, 224, device=dev); y = torch.randint(0, 1000, (bs,), device=dev)
    F.cross_entropy(ref(x), y).backward()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        la = F.cross_entropy(amp(x), y)
    la.backward()
    Fn.cross_entropy(mine(x), y).backward()
    names = ["fc.weight", "fc.bias", "layer4.2.bn3.weight", "layer4.2.conv3.weight", "layer4.2.conv2.weight", "layer4.2.conv1.weight",
             "layer4.0.downsample.0.weight", "layer4.0.conv2.weight", "layer3.5.conv3.weight", "layer3.0.conv2.weight", "layer2.0.conv1.weight",
             "layer1.0.conv1.weight", "layer1.0.conv2.weight", "bn1.weight", "conv1.weight"]
    pm, pr, pa = dict(mine.named_parameters()), dict(ref.named_parameters()), dict(amp.named_parameters())
    print(f"batch {bs}: name | cos(mine,fp32) | cos(autocast,fp32) | |g|mine/|g|fp32")
    for n in names:
        g = pm[n].grad
        if g.dim() == 4:
            g = g.permute(0, 3, 1, 2)
        a, b, c = g.flatten().float(), pr[n].grad.flatten().float(), pa[n].grad.flatten().float()
        print(f"  {n:32s} {F.cosine_similarity(a, b, dim=0).item():.4f}  {F.cosine_similarity(c, b, dim=0).item():.4f}  {(a.norm() / b.norm()).item():.3f}")
