import sys, os, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pytorch_distributed_training_tutorials_b200.models.resnet_mp import PipelineParallelResNet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn

def cosv(a, b):
    a, b = a.flatten().double(), b.flatten().double()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-30))

devs = ["cuda:0", "cuda:1" if torch.cuda.device_count() > 1 else "cuda:0"]
for chunks in (1, 2):
    torch.manual_seed(0)
    m = PipelineParallelResNet50(split_size=16 // chunks, num_classes=16, dev0=devs[0], dev1=devs[1], split_after="layer1")
    r = copy.deepcopy(m); r2 = copy.deepcopy(m)
    x = torch.randn(16, 3, 64, 64, device=devs[0]); y = torch.randint(0, 16, (16,), device=devs[1])
    pipe = m.graphed(Fn.cross_entropy, x, y)
    loss = pipe.train_step(x, y)
    for d in devs: torch.cuda.synchronize(d)
    # eager, slot by slot
    tot = 0
    for i, (xi, yi) in enumerate(zip(x.chunk(chunks), y.chunk(chunks))):
        a0 = r.stage0(xi)
        a1 = a0.detach().to(devs[1]).requires_grad_(True)
        li = Fn.cross_entropy(r.stage1(a1), yi) / chunks
        li.backward()
        a0.backward(a1.grad.to(devs[0]))
        tot += float(li)
        print(f"chunks {chunks} slot {i}: stage0 out cos {cosv(pipe.out[0][i], a0):.5f}  loss_i {float(pipe.out[1][i]):.5f} vs {float(li):.5f}"
              f"  gin cos {cosv(pipe.gin[1][i], a1.grad):.5f}")
    # eager twice: noise yardstick
    for xi, yi in zip(x.chunk(chunks), y.chunk(chunks)):
        (Fn.cross_entropy(r2(xi), yi) / chunks).backward()
    names = [n for n, _ in m.named_parameters()]
    cg = [cosv(p.grad, q.grad) for p, q in zip(m.parameters(), r.parameters())]
    ce = [cosv(p.grad, q.grad) for p, q in zip(r2.parameters(), r.parameters())]
    print(f"chunks {chunks}: loss {float(loss):.5f} vs {tot:.5f}; grad cos graph-vs-eager min {min(cg):.4f} mean {sum(cg)/len(cg):.4f};"
          f" eager-vs-eager min {min(ce):.4f} mean {sum(ce)/len(ce):.4f}")
    worst = sorted(zip(cg, names))[:5]
    print("  worst:", worst)
    print("  fc.weight", cosv(m.fc.weight.grad, r.fc.weight.grad), "layer4.2.conv3", cosv(m.layer4[2].conv3.weight.grad, r.layer4[2].conv3.weight.grad),
          "conv1", cosv(m.conv1.weight.grad, r.conv1.weight.grad))
