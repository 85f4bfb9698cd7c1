"""Run-to-run reproducibility of the kernels and of a whole ResNet-50 step.

Three questions, each answered with numbers (JSON on stdout, one object per line):
 1. kernel level: does the same launch on the same inputs give the same bits twice?  (outputs that are
    produced without atomics must; fp32 outputs that are summed with atomics may differ in the last ulp
    unless B200DT_DETERMINISTIC=1)
 2. model level: where do two runs of the same model on the same batch first differ, and by how much?
 3. yardstick: how much does torchvision's own ResNet-50 (library kernels, bf16 autocast) move when the input
    is perturbed by 1e-6 relative?  That is the conditioning of the problem, not a property of any kernel.
"""
import argparse
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from pytorch_distributed_training_tutorials_b200.models.resnet import resnet50
from pytorch_distributed_training_tutorials_b200.ops import functional as Fn
from pytorch_distributed_training_tutorials_b200.ops import native as NV
from pytorch_distributed_training_tutorials_b200.ops import nn as bnn

dev = torch.device("cuda:0")


def cos(a, b):
    a, b = a.flatten().double(), b.flatten().double()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-300))


def diff(a, b):
    a32, b32 = a.float(), b.float()
    ne = int((a32 != b32).sum())
    return {"neq": ne, "frac": ne / max(1, a.numel()),
            "rel": float((a32 - b32).norm() / (b32.norm() + 1e-30))}


def kernel_level(nb):
    shapes = [(56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (56, 256, 64, 1, 1), (56, 256, 128, 1, 1), (56, 128, 128, 3, 2),
              (28, 128, 512, 1, 1), (28, 512, 128, 1, 1), (28, 128, 128, 3, 1), (56, 256, 512, 1, 2), (14, 256, 256, 3, 1),
              (14, 1024, 256, 1, 1), (7, 512, 512, 3, 1), (7, 512, 2048, 1, 1)]
    for (H, Cin, Cout, R, s) in shapes:
        torch.manual_seed(H * 1000 + Cin + Cout + R)
        x = torch.randn(nb, H, H, Cin, device=dev).to(torch.bfloat16)
        w = (torch.randn(Cout, R, R, Cin, device=dev) * (2.0 / (Cin * R * R)) ** 0.5).to(torch.bfloat16)
        dy = torch.randn(nb, H // s, H // s, Cout, device=dev).to(torch.bfloat16)
        skip = torch.randn(nb, H, H, Cin, device=dev).to(torch.bfloat16)
        out = {"kind": "kernel", "shape": [nb, H, Cin, Cout, R, s]}
        ys, sums, dxs, accs, dws = [], [], [], [], []
        for rep in range(3):
            sb = NV.bn_sums_buffer(Cout, dev)
            sb.fill_(float("nan"))      # the kernel must not depend on what the scratch held
            ys.append(NV.conv2d_fprop(x, w, stride=s, bn_sums=sb))
            sums.append(sb[:2 * Cout].clone())
            dxs.append(NV.conv2d_dgrad(dy, w, x.shape, R, stride=s))
            accs.append(NV.conv2d_dgrad(dy, w, x.shape, R, stride=s, accumulate_into=skip.clone()))
            dwo = torch.full((Cout, R, R, Cin), float("nan"), device=dev)
            dws.append(NV.conv2d_wgrad(x, dy, R, stride=s, out=dwo))
            # disturb the allocator / L2 between repetitions
            junk = torch.randn(1 << 22, device=dev)
            del junk
        torch.cuda.synchronize()
        out["y"] = max(diff(ys[i], ys[0])["neq"] for i in (1, 2))
        out["dx"] = max(diff(dxs[i], dxs[0])["neq"] for i in (1, 2))
        out["dx_acc"] = max(diff(accs[i], accs[0])["neq"] for i in (1, 2))
        out["sums_rel"] = max(diff(sums[i], sums[0])["rel"] for i in (1, 2))
        out["sums_neq"] = max(diff(sums[i], sums[0])["neq"] for i in (1, 2))
        out["dw_rel"] = max(diff(dws[i], dws[0])["rel"] for i in (1, 2))
        out["dw_neq"] = max(diff(dws[i], dws[0])["neq"] for i in (1, 2))
        # fp32 reference of the statistics and of dgrad_acc = skip + dgrad
        yf = ys[0].float()
        ref_s = torch.cat([yf.sum((0, 1, 2)), (yf * yf).sum((0, 1, 2))])
        out["sums_vs_ref"] = float((sums[0] - ref_s).norm() / ref_s.norm())
        out["acc_vs_ref"] = float((accs[0].float() - (skip.float() + dxs[0].float())).norm() / accs[0].float().norm())
        # BN forward/backward on this output
        g = torch.rand(Cout, device=dev) + 0.5
        b = torch.randn(Cout, device=dev) * 0.1
        res = torch.randn_like(ys[0])
        bo = []
        for rep in range(2):
            rm, rv = torch.zeros(Cout, device=dev), torch.ones(Cout, device=dev)
            x2 = ys[0].reshape(-1, Cout)
            mask = torch.empty(x2.numel() // 8, device=dev, dtype=torch.uint8)
            yb, mean, invstd = NV.bn_apply(x2, res.reshape(-1, Cout), sums_full(sums[0], Cout), g, b, rm, rv, 1e-5, 0.1, True, None, mask)
            dxb, dg, db, dres = NV.bn_backward(dy.reshape(-1, Cout), None, x2, mean, invstd, g, True, True, beta=b, relu_mask=mask)
            bo.append((yb, dxb, dg, db, dres))
        torch.cuda.synchronize()
        out["bn_y"] = diff(bo[1][0], bo[0][0])["neq"]
        out["bn_dx"] = diff(bo[1][1], bo[0][1])["neq"]
        out["bn_dgamma_rel"] = diff(bo[1][2], bo[0][2])["rel"]
        out["bn_dres"] = diff(bo[1][4], bo[0][4])["neq"]
        print(json.dumps(out), flush=True)


def sums_full(s2c, C):
    full = NV.bn_sums_buffer(C, dev)
    full.zero_()
    full[:2 * C] = s2c
    return full


def run_model(m, x, y, loss_fn, outs):
    hooks = []
    for name, mod in m.named_modules():
        if isinstance(mod, (bnn.BatchNorm2d, bnn.MaxPool3x3s2, bnn.GlobalAvgPool, bnn.Linear)):
            hooks.append(mod.register_forward_hook(lambda _m, _i, o, name=name: outs.append((name, o.detach().clone()))))
    m.zero_grad(set_to_none=True)
    loss = loss_fn(m(x), y)
    loss.backward()
    for h in hooks:
        h.remove()
    return float(loss)


def model_level(nb, hw, classes, zero_init):
    torch.manual_seed(0)
    m = resnet50(num_classes=classes, zero_init_residual=zero_init).to(dev).train()
    m2 = copy.deepcopy(m)
    x = torch.randn(nb, 3, hw, hw, device=dev)
    y = torch.randint(0, classes, (nb,), device=dev)
    o1, o2, o3 = [], [], []
    l1 = run_model(m, x, y, Fn.cross_entropy, o1)
    g1 = {n: p.grad.clone() for n, p in m.named_parameters()}
    l2 = run_model(m, x, y, Fn.cross_entropy, o2)          # same module object, second run
    g2 = {n: p.grad.clone() for n, p in m.named_parameters()}
    l3 = run_model(m2, x, y, Fn.cross_entropy, o3)         # deep copy: different addresses
    g3 = {n: p.grad.clone() for n, p in m2.named_parameters()}
    rep = {"kind": "model", "nb": nb, "hw": hw, "classes": classes, "zero_init": zero_init,
           "det": os.environ.get("B200DT_DETERMINISTIC", "0"), "loss": [l1, l2, l3]}
    first = None
    worst = 0.0
    for (n, a), (_, b), (_, c) in zip(o1, o2, o3):
        d = max(diff(a, b)["rel"], diff(a, c)["rel"])
        worst = max(worst, d)
        if first is None and d > 0:
            first = (n, diff(a, b), diff(a, c))
    rep["fwd_first_diff"] = first
    rep["fwd_worst_rel"] = worst
    rep["fwd_last_rel"] = diff(o1[-1][1], o2[-1][1])["rel"]
    cs = {n: min(cos(g1[n], g2[n]), cos(g1[n], g3[n])) for n in g1}
    neq = {n: max(diff(g1[n], g2[n])["neq"], diff(g1[n], g3[n])["neq"]) for n in g1}
    rep["grad_cos_mean"] = sum(cs.values()) / len(cs)
    rep["grad_cos_min"] = min(cs.items(), key=lambda kv: kv[1])
    rep["grad_bit_identical_params"] = sum(1 for v in neq.values() if v == 0)
    rep["grad_params"] = len(neq)
    for n in ("fc.weight", "layer4.2.conv3.weight", "layer3.0.conv2.weight", "layer2.0.conv1.weight", "layer1.0.conv1.weight", "conv1.weight"):
        rep["cos:" + n] = cs[n]
    print(json.dumps(rep), flush=True)


def yardstick(nb, hw, classes, zero_init, layers=(3, 4, 6, 3)):
    import torchvision
    from torchvision.models.resnet import Bottleneck, ResNet as TVResNet
    from pytorch_distributed_training_tutorials_b200.models.resnet import ResNet
    torch.manual_seed(0)
    ref = TVResNet(Bottleneck, list(layers), num_classes=classes, zero_init_residual=zero_init).to(dev).train()
    x = torch.randn(nb, 3, hw, hw, device=dev)
    y = torch.randint(0, classes, (nb,), device=dev)
    xp = x * (1 + 1e-6 * torch.randn_like(x))
    res = {"kind": "yardstick", "nb": nb, "hw": hw, "classes": classes, "zero_init": zero_init, "layers": list(layers)}
    for tag, amp in (("fp32", False), ("bf16", True)):
        gs = []
        for inp in (x, xp):
            ref.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                l = F.cross_entropy(ref(inp), y)
            l.backward()
            gs.append({n: p.grad.clone() for n, p in ref.named_parameters()})
        cs = [cos(gs[0][n], gs[1][n]) for n in gs[0]]
        res[tag + "_cos_mean_under_1e-6_input_noise"] = sum(cs) / len(cs)
        res[tag + "_cos_conv1"] = cos(gs[0]["conv1.weight"], gs[1]["conv1.weight"])
        if tag == "fp32":
            g32 = gs[0]
        else:
            cs = [cos(gs[0][n], g32[n]) for n in g32]
            res["bf16_vs_fp32_cos_mean"] = sum(cs) / len(cs)
            res["bf16_vs_fp32_cos_conv1"] = cos(gs[0]["conv1.weight"], g32["conv1.weight"])
    # ours vs fp32
    mine = ResNet(list(layers), num_classes=classes, zero_init_residual=zero_init).to(dev).train()
    mine.load_state_dict(ref.state_dict())
    Fn.cross_entropy(mine(x), y).backward()
    pm = dict(mine.named_parameters())
    cs = {}
    for n in g32:
        g = pm[n].grad
        g = g.permute(0, 3, 1, 2) if g.dim() == 4 else g
        cs[n] = cos(g, g32[n])
    res["ours_vs_fp32_cos_mean"] = sum(cs.values()) / len(cs)
    res["ours_vs_fp32_cos_min"] = min(cs.items(), key=lambda kv: kv[1])
    res["ours_vs_fp32_cos_conv1"] = cs["conv1.weight"]
    res["ours_vs_fp32_cos_all"] = {k: round(v, 5) for k, v in cs.items()} if len(cs) < 80 else None
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="kernel,model,yardstick")
    ap.add_argument("--det", type=int, default=-1, help="1: deterministic reductions, 0: default, -1: keep the env setting")
    a = ap.parse_args()
    if a.det >= 0:
        from pytorch_distributed_training_tutorials_b200 import _lib
        _lib.set_deterministic(bool(a.det))
        os.environ["B200DT_DETERMINISTIC"] = str(a.det)
    what = a.what.split(",")
    if "kernel" in what:
        kernel_level(8)
    if "model" in what:
        model_level(8, 128, 16, False)
        model_level(64, 128, 16, True)
        model_level(32, 224, 1000, False)
    if "yardstick" in what:
        yardstick(8, 128, 16, False)
        yardstick(64, 128, 16, True)
        yardstick(32, 224, 1000, False)
    if "shallow" in what:
        yardstick(64, 128, 1000, False, layers=(1, 1, 1, 1))
        yardstick(64, 224, 1000, False, layers=(1, 1, 1, 1))
