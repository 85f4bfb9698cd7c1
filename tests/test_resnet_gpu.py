"""ResNet-50 on the sm_100a kernels vs torchvision (fp32 reference), forward and backward."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = a.float(); b = b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def test_resnet50_matches_torchvision():
    import torchvision
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    ref = torchvision.models.resnet50(num_classes=1000).to(dev).train()
    mine = resnet50(num_classes=1000).to(dev).train()
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(16, 3, 224, 224, device=dev)
    y = torch.randint(0, 1000, (16,), device=dev)
    out_ref = ref(x)
    F.cross_entropy(out_ref, y).backward()
    out = mine(x)
    assert out.shape == (16, 1000) and out.dtype == torch.float32
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    loss = Fn.cross_entropy(out, y)
    loss.backward()
    assert rel(out, out_ref) < 5e-2
    # BN running statistics follow the same update rule
    assert rel(mine.bn1.running_var, ref.bn1.running_var) < 2e-2
    assert rel(mine.layer4[2].bn3.running_mean, ref.layer4[2].bn3.running_mean) < 5e-2
    # gradients: bf16 activations through 50 layers; compare direction and scale
    pairs = [(mine.fc.weight.grad, ref.fc.weight.grad),
             (mine.layer4[2].conv3.weight.grad.permute(0, 3, 1, 2), ref.layer4[2].conv3.weight.grad),
             (mine.layer3[0].conv2.weight.grad.permute(0, 3, 1, 2), ref.layer3[0].conv2.weight.grad),
             (mine.layer1[0].conv1.weight.grad.permute(0, 3, 1, 2), ref.layer1[0].conv1.weight.grad),
             (mine.conv1.weight.grad.permute(0, 3, 1, 2), ref.conv1.weight.grad),
             (mine.bn1.weight.grad, ref.bn1.weight.grad)]
    for i, (a, b) in enumerate(pairs):
        cos = F.cosine_similarity(a.flatten().float(), b.flatten().float(), dim=0).item()
        assert cos > 0.95, (i, cos, rel(a, b))


def test_resnet50_trains_and_checkpoint_roundtrip(tmp_path):
    import torchvision
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    torch.manual_seed(1)
    dev = torch.device("cuda:0")
    m = resnet50(num_classes=10).to(dev)
    opt = FusedSGD(m.parameters(), lr=0.05)
    x = torch.randn(32, 3, 64, 64, device=dev)
    y = torch.randint(0, 10, (32,), device=dev)
    losses = []
    for _ in range(8):
        opt.zero_grad()
        loss = Fn.cross_entropy(m(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0], losses          # it overfits one batch
    torch.save(m.state_dict(), tmp_path / "m.pt")
    tv = torchvision.models.resnet50(num_classes=10)
    tv.load_state_dict(torch.load(tmp_path / "m.pt"))  # our checkpoints load into plain torchvision
