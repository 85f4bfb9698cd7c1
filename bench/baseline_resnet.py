"""Quick single-GPU measurement of the reference stack (torchvision ResNet-50 + autocast bf16 + SGD)."""
import sys, time, json, os
import torch, torchvision
import torch.nn.functional as F

bs = int(sys.argv[1]) if len(sys.argv) > 1 else 256
res = {}
for cl in (False, True):
    torch.manual_seed(0)
    m = torchvision.models.resnet50(num_classes=1000).cuda()
    if cl:
        m = m.to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(m.parameters(), lr=1e-3)
    x = torch.randn(bs, 3, 224, 224, device="cuda")
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (bs,), device="cuda")

    def step():
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = m(x)
            loss = F.cross_entropy(out, y)
        loss.backward()
        opt.step()

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 10
    for _ in range(n):
        step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    res["channels_last" if cl else "nchw"] = {"ms_per_step": ms, "img_s": bs / ms * 1e3}
    print(("channels_last" if cl else "nchw"), f"{ms:.2f} ms/step  {bs / ms * 1e3:.0f} img/s", flush=True)
    del m, opt
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(f"gpurun_out/baseline_resnet_bs{bs}.json", "w"))
