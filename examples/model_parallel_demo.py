"""Model parallelism (counterpart of /root/reference/03.model_parallel.ipynb).

    python examples/model_parallel_demo.py toy        # Linear(10000,10) on cuda:0 -> Linear(10,5) on cuda:1
    python examples/model_parallel_demo.py resnet     # ResNet-50 split after layer2, naive vs pipelined
    python examples/model_parallel_demo.py place      # LLaMA-shaped stack auto-placed over the GPUs

Every mode accepts ``--cpu`` to walk through the same code with CPU placeholders (``resnet --cpu`` runs the naive and
the eager pipelined split on the modules' plain-torch reference path; the CUDA-graph replay needs GPUs).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn

import pytorch_distributed_training_tutorials_b200 as b2


def toy(cpu: bool) -> None:
    from pytorch_distributed_training_tutorials_b200.models.toy import ToyModel
    d0, d1 = ("cpu", "cpu") if cpu else ("cuda:0", "cuda:1")
    model = ToyModel(d0, d1)
    print(next(model.net1.parameters()).device, next(model.net2.parameters()).device)
    loss_fn = nn.MSELoss()
    optimizer = (torch.optim.SGD if cpu else b2.FusedSGD)(model.parameters(), lr=0.001)
    optimizer.zero_grad()
    outputs = model(torch.randn(20, 10000))
    labels = torch.randn(20, 5).to(d1)
    loss_fn(outputs, labels).backward()
    optimizer.step()
    print("one optimizer step over parameters on two devices: ok, loss device", outputs.device)


def resnet(cpu: bool = False) -> None:
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
    from pytorch_distributed_training_tutorials_b200.parallel.model_parallel import model_size
    dev = dict(dev0="cpu", dev1="cpu") if cpu else {}
    print("resnet50 parameters:", model_size(resnet50(num_classes=1000)))
    mp = ModelParallelResNet50(num_classes=1000, **dev)
    print("split resnet50 parameters:", model_size(mp), "| stage devices:", mp.conv1.weight.device, mp.fc.weight.device)
    pp = PipelineParallelResNet50(split_size=4 if cpu else 20, num_classes=1000, split_after="layer2.1", **dev)
    x = torch.randn(8, 3, 64, 64) if cpu else torch.randn(120, 3, 128, 128)
    print("naive:", tuple(mp(x).shape), "pipelined (cut after layer2.1):", tuple(pp(x).shape))
    if cpu:
        return
    # the fast pipelined form: every (stage, micro-batch) is captured once and replayed from CUDA graphs
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    gp = PipelineParallelResNet50(split_size=60, num_classes=1000, split_after="layer1")
    opt = FusedSGD(gp.parameters(), lr=1e-3)
    labels = torch.zeros(120, 1000).scatter_(1, torch.randint(0, 1000, (120, 1)), 1)
    pipe = gp.graphed(Fn.mse_loss, x.to(gp.dev0), labels.to(gp.dev1))
    for _ in range(3):
        pipe.zero_grad()
        loss = pipe.train_step(x, labels)
        opt.step()
    print(f"graph-replayed pipeline ({pipe.chunks} micro-batches, {pipe.launches_per_step} graph launches per step): "
          f"loss {float(loss):.4f}")
    print("timings: python bench/mp_bench.py ; python bench/graphed_pipeline_bench.py")


def place(cpu: bool) -> None:
    from pytorch_distributed_training_tutorials_b200.models.llama_like import LlamaLikeForCausalLM
    from pytorch_distributed_training_tutorials_b200.parallel.place import auto_place, dispatch, placement_table
    n = 4 if cpu else max(torch.cuda.device_count(), 1)
    model = LlamaLikeForCausalLM(vocab=1024, dim=256, layers=32, heads=8, hidden=688)
    devices = [torch.device("cpu")] * n if cpu else [torch.device("cuda", i) for i in range(n)]
    placement = auto_place(model, devices, blocks=model.placement_blocks())
    dispatch(model, placement)
    for i, name, dtype, dev in placement_table(model):
        print(i, name, dtype, dev)
    out = model(torch.randint(0, 1024, (2, 16), device=devices[0]))
    print("forward through", n, "placements ok:", tuple(out.shape), out.device)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["toy", "resnet", "place"])
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()
    {"toy": lambda: toy(a.cpu), "resnet": lambda: resnet(a.cpu), "place": lambda: place(a.cpu)}[a.what]()
