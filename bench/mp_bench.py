"""ResNet-50 on one GPU vs split across two (naive and pipelined): BASELINE.json config 4.

Workload of the reference (/root/reference/03.model_parallel.ipynb:969-992): 3 batches of 120 images
3x128x128 created on the host, one-hot labels, MSELoss, SGD lr 1e-3.  The reference times
``train(model)`` with host ``timeit`` (mean of 10, no synchronize) and draws a bar chart; here every
variant is timed on the device (CUDA events around the whole ``train``, synchronised) AND on the host,
and the chart is replaced by JSON + a markdown table (no graphics stack on this box).

Variants: this library (single / naive MP / pipelined MP with a split-size sweep) and the reference
stack: torchvision ResNet-50 and the notebook's own ``ModelParallelResNet50`` + ``train`` cells,
executed unmodified from baseline/_ref/03.model_parallel.ipynb.
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn
import torch.optim as optim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
num_batches, batch_size, image_w, image_h, num_classes = 3, 120, 128, 128, 1000


def train_ours(model, in_dev, out_dev, opt):
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    model.train(True)
    idx = torch.LongTensor(batch_size).random_(0, num_classes).view(batch_size, 1)
    for _ in range(num_batches):
        inputs = torch.randn(batch_size, 3, image_w, image_h)
        labels = torch.zeros(batch_size, num_classes).scatter_(1, idx, 1)
        opt.zero_grad()
        outputs = model(inputs.to(in_dev))
        Fn.mse_loss(outputs, labels.to(outputs.device)).backward()
        opt.step()


def timed(fn, repeats=10):
    fn()                                       # warm-up (the reference warms MP only: nb03:1004)
    dev_ms, host_s = [], []
    for _ in range(repeats):
        for d in range(torch.cuda.device_count()):
            torch.cuda.synchronize(d)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        with torch.cuda.device(0):
            e0.record()
        fn()
        host_s.append(time.perf_counter() - t0)        # like timeit: until the last op is enqueued
        for d in range(torch.cuda.device_count()):
            torch.cuda.synchronize(d)
        with torch.cuda.device(0):
            e1.record(); e1.synchronize()
        dev_ms.append(e0.elapsed_time(e1))
    t = torch.tensor(dev_ms)
    return {"device_ms_mean": float(t.mean()), "device_ms_std": float(t.std()), "host_s_mean": sum(host_s) / len(host_s),
            "img_per_s": num_batches * batch_size / (float(t.mean()) / 1e3)}


def reference_cells():
    path = os.path.join(ROOT, "baseline", "_ref", "03.model_parallel.ipynb")
    if not os.path.exists(path):
        return None
    nb = json.load(open(path))
    ns = {"torch": torch, "nn": nn, "optim": optim}
    exec("from torchvision.models.resnet import ResNet, Bottleneck\nnum_classes = 1000", ns)
    for cell in nb["cells"]:
        src = "".join(cell.get("source", []))
        if cell.get("cell_type") == "code" and ("class ModelParallelResNet50" in src or "def train(model)" in src):
            exec(src, ns)                      # the notebook's own code, unmodified
    return ns


def gpu_bound(res):
    """Second workload: the same models with batch 1024 at 224x224 and inputs already on the GPU, so a
    micro-batch of 256 keeps a B200 busy for longer than the host needs to issue it (the notebook's
    120 x 128^2 batches are bound by the host-side randn() and launch overhead on this hardware)."""
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    B, S, steps = 1024, 224, 3
    x = torch.randn(B, 3, S, S, device="cuda:0")
    lab = torch.zeros(B, num_classes).scatter_(1, torch.randint(0, num_classes, (B, 1)), 1)

    def run(model, opt, out_dev, loss):
        y = lab.to(out_dev)
        def f():
            for _ in range(steps):
                opt.zero_grad()
                loss(model(x), y).backward()
                opt.step()
        r = timed(f, repeats=5)
        r["img_per_s"] = steps * B / (r["device_ms_mean"] / 1e3)
        return r
    m = resnet50(num_classes=num_classes).to("cuda:0"); o = FusedSGD(m.parameters(), lr=1e-3)
    res["gpu_bound/ours_single_gpu"] = run(m, o, "cuda:0", Fn.mse_loss); del m, o
    if torch.cuda.device_count() >= 2:
        m = ModelParallelResNet50(num_classes=num_classes); o = FusedSGD(m.parameters(), lr=1e-3)
        res["gpu_bound/ours_naive_mp"] = run(m, o, "cuda:1", Fn.mse_loss); del m, o
        for cut in ("layer2", "layer1"):
            for split in (256, 512):
                m = PipelineParallelResNet50(split_size=split, num_classes=num_classes, split_after=cut); o = FusedSGD(m.parameters(), lr=1e-3)
                res[f"gpu_bound/ours_pipelined_mp_cut_{cut}_split{split}"] = run(m, o, "cuda:1", Fn.mse_loss); del m, o
        m = ModelParallelResNet50(num_classes=num_classes, split_after="layer1"); o = FusedSGD(m.parameters(), lr=1e-3)
        res["gpu_bound/ours_naive_mp_cut_layer1"] = run(m, o, "cuda:1", Fn.mse_loss); del m, o
    ns = reference_cells()
    if ns is not None:
        import torchvision.models as models
        mse = nn.MSELoss()
        rn = models.resnet50(num_classes=num_classes).to("cuda:0"); ro = optim.SGD(rn.parameters(), lr=1e-3)
        res["gpu_bound/reference_single_gpu_fp32"] = run(rn, ro, "cuda:0", mse); del rn, ro
        if torch.cuda.device_count() >= 2:
            mp = ns["ModelParallelResNet50"](); ro = optim.SGD(mp.parameters(), lr=1e-3)
            res["gpu_bound/reference_naive_mp_fp32"] = run(mp, ro, "cuda:1", mse); del mp, ro


def main():
    from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
    from pytorch_distributed_training_tutorials_b200.models.resnet_mp import ModelParallelResNet50, PipelineParallelResNet50
    from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
    res = {}
    m = resnet50(num_classes=num_classes).to("cuda:0"); o = FusedSGD(m.parameters(), lr=1e-3)
    res["ours_single_gpu"] = timed(lambda: train_ours(m, "cuda:0", "cuda:0", o))
    del m, o
    if torch.cuda.device_count() >= 2:
        m = ModelParallelResNet50(num_classes=num_classes); o = FusedSGD(m.parameters(), lr=1e-3)
        res["ours_naive_mp"] = timed(lambda: train_ours(m, "cuda:0", "cuda:1", o))
        del m, o
        for split in (20, 40, 60):
            m = PipelineParallelResNet50(split_size=split, num_classes=num_classes); o = FusedSGD(m.parameters(), lr=1e-3)
            res[f"ours_pipelined_mp_split{split}"] = timed(lambda: train_ours(m, "cuda:0", "cuda:1", o))
            del m, o
        # CUDA-graph replayed pipeline on the same workload (inputs still created on the host per batch)
        from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
        idx = torch.LongTensor(batch_size).random_(0, num_classes).view(batch_size, 1)
        for split in (30, 60):
            m = PipelineParallelResNet50(split_size=split, num_classes=num_classes, split_after="layer1")
            o = FusedSGD(m.parameters(), lr=1e-3)
            ex_x = torch.randn(batch_size, 3, image_w, image_h, device="cuda:0")
            ex_y = torch.zeros(batch_size, num_classes, device="cuda:1")
            pipe = m.graphed(Fn.mse_loss, ex_x, ex_y)

            def train_graphed():
                for _ in range(num_batches):
                    inputs = torch.randn(batch_size, 3, image_w, image_h)
                    labels = torch.zeros(batch_size, num_classes).scatter_(1, idx, 1)
                    pipe.zero_grad()
                    pipe.train_step(inputs, labels)
                    o.step()
            res[f"ours_pipelined_graphed_cut_layer1_split{split}"] = timed(train_graphed)
            del pipe, m, o
        # the naive split and the single GPU, graph-replayed (same host-side input creation per batch)
        m = ModelParallelResNet50(num_classes=num_classes); o = FusedSGD(m.parameters(), lr=1e-3)
        ex_x = torch.randn(batch_size, 3, image_w, image_h, device="cuda:0")
        ex_y = torch.zeros(batch_size, num_classes, device="cuda:1")
        pipe = m.graphed(Fn.mse_loss, ex_x, ex_y)

        def train_naive_graphed():
            for _ in range(num_batches):
                inputs = torch.randn(batch_size, 3, image_w, image_h)
                labels = torch.zeros(batch_size, num_classes).scatter_(1, idx, 1)
                pipe.zero_grad()
                pipe.train_step(inputs, labels)
                o.step()
        res["ours_naive_mp_graphed"] = timed(train_naive_graphed)
        del pipe, m, o
    from pytorch_distributed_training_tutorials_b200.utils.graph import GraphedTrainStep
    from pytorch_distributed_training_tutorials_b200.ops import functional as Fn2
    m = resnet50(num_classes=num_classes).to("cuda:0"); o = FusedSGD(m.parameters(), lr=1e-3)
    g = GraphedTrainStep(m, Fn2.mse_loss, o, torch.randn(batch_size, 3, image_w, image_h, device="cuda:0"),
                         torch.zeros(batch_size, num_classes, device="cuda:0"))
    idx1 = torch.LongTensor(batch_size).random_(0, num_classes).view(batch_size, 1)

    def train_single_graphed():
        for _ in range(num_batches):
            inputs = torch.randn(batch_size, 3, image_w, image_h)
            labels = torch.zeros(batch_size, num_classes).scatter_(1, idx1, 1)
            g(inputs.to("cuda:0"), labels.to("cuda:0"))
    res["ours_single_gpu_graphed"] = timed(train_single_graphed)
    del g, m, o
    ns = reference_cells()
    if ns is not None:
        import torchvision.models as models
        rn = models.resnet50(num_classes=num_classes).to("cuda:0")
        res["reference_single_gpu"] = timed(lambda: ns["train"](rn))
        del rn
        if torch.cuda.device_count() >= 2:
            mp = ns["ModelParallelResNet50"]()
            res["reference_naive_mp"] = timed(lambda: ns["train"](mp))
    gpu_bound(res)
    print("| variant | device ms / train() (mean +- std) | host s (timeit-style) | img/s |\n|---|---|---|---|")
    for k, v in res.items():
        print(f"| {k} | {v['device_ms_mean']:.1f} +- {v['device_ms_std']:.1f} | {v['host_s_mean']:.3f} | {v['img_per_s']:.0f} |")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"workload": "3 x 120 x 3x128x128, MSELoss, SGD 1e-3", "published_reference": {"naive_mp_s": 0.346, "single_gpu_s": 0.374},
               "results": res}, open("gpurun_out/mp_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
