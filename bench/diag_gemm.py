"""GPU diagnostic for the tcgen05 GEMM / implicit-GEMM kernel: prints an error map per case so that
one gpurun call tells which operand path (TMA box, swizzle, descriptor, epilogue) is off."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from pytorch_distributed_training_tutorials_b200.ops import native as NV

dev = torch.device("cuda:0")
results = []


def emap(out, ref, bs=32):
    out = out.float(); ref = ref.float()
    M, N = out.shape
    rows = []
    for i in range(0, min(M, 256), bs):
        row = []
        for j in range(0, min(N, 256), bs):
            o = out[i:i + bs, j:j + bs]; r = ref[i:i + bs, j:j + bs]
            row.append(float((o - r).abs().max() / (r.abs().max() + 1e-6)))
        rows.append(" ".join(f"{v:7.3f}" for v in row))
    return "\n".join(rows)


def report(name, out, ref, tol):
    err = ((out.float() - ref.float()).norm() / (ref.float().norm() + 1e-12)).item()
    ok = err < tol
    print(f"[{'OK ' if ok else 'BAD'}] {name}: rel_err={err:.3e} out_absmax={out.float().abs().max().item():.3e} "
          f"ref_absmax={ref.float().abs().max().item():.3e} nan={bool(torch.isnan(out.float()).any())}", flush=True)
    if not ok:
        o2 = out.float().reshape(-1, out.shape[-1]); r2 = ref.float().reshape(-1, ref.shape[-1])
        print(emap(o2, r2), flush=True)
    results.append({"name": name, "rel_err": err, "ok": ok})
    return ok


def run(name, fn):
    try:
        fn()
        torch.cuda.synchronize()
    except Exception as e:  # noqa
        print(f"[EXC] {name}: {type(e).__name__}: {e}", flush=True)
        results.append({"name": name, "ok": False, "exc": str(e)})


def gemm_case(M, N, K):
    torch.manual_seed(0)
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    out = NV.gemm_tn(a, b, out_dtype=torch.float32)
    report(f"gemm_tn {M}x{N}x{K}", out, a.float() @ b.float().t(), 2e-3)


def gemm_struct():
    # structured operands: isolates row / column / k mapping
    M, N, K = 128, 64, 64
    a = torch.zeros(M, K, device=dev); a[:, 3] = torch.arange(M, device=dev).float() % 7 + 1
    b = torch.zeros(N, K, device=dev); b[:, 3] = torch.arange(N, device=dev).float() % 5 + 1
    out = NV.gemm_tn(a.bfloat16(), b.bfloat16(), out_dtype=torch.float32)
    report("gemm_struct k=3", out, a @ b.t(), 1e-3)
    a = torch.zeros(M, K, device=dev); b = torch.zeros(N, K, device=dev)
    a[:, 40] = 1; b[:, 40] = torch.arange(N, device=dev).float() + 1
    out = NV.gemm_tn(a.bfloat16(), b.bfloat16(), out_dtype=torch.float32)
    report("gemm_struct k=40", out, a @ b.t(), 1e-3)


def nt_case(R, M, N):
    torch.manual_seed(0)
    a = torch.randn(R, M, device=dev).bfloat16(); b = torch.randn(R, N, device=dev).bfloat16()
    out = NV.gemm_nt_reduce(a, b)
    report(f"gemm_nt_reduce R={R} {M}x{N}", out, a.float().t() @ b.float(), 2e-3)


def conv_case(NB, H, W, Cin, Cout, R, s):
    torch.manual_seed(1)
    x = torch.randn(NB, H, W, Cin, device=dev).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=dev) / (R * Cin ** 0.5)).bfloat16()
    dy = torch.randn(NB, H // s, W // s, Cout, device=dev).bfloat16()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)
    yref = F.conv2d(xr, wr, stride=s, padding=R // 2)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    tag = f"NB{NB} {H}x{W} {Cin}->{Cout} R{R} s{s}"
    run("fprop " + tag, lambda: report("conv_fprop " + tag, NV.conv2d_fprop(x, w, stride=s), yref.permute(0, 2, 3, 1), 1e-2))
    run("wgrad " + tag, lambda: report("conv_wgrad " + tag, NV.conv2d_wgrad(x, dy, R, stride=s), wr.grad.permute(0, 2, 3, 1), 5e-3))
    run("dgrad " + tag, lambda: report("conv_dgrad " + tag, NV.conv2d_dgrad(dy, w, x.shape, R, stride=s), xr.grad.permute(0, 2, 3, 1), 1e-2))


run("struct", gemm_struct)
for c in [(128, 64, 64), (128, 128, 64), (128, 256, 64), (128, 64, 256), (256, 128, 512), (1000, 1000, 2048), (300, 72, 200), (8192, 2048, 1024)]:
    run(f"gemm {c}", lambda c=c: gemm_case(*c))
for c in [(64, 64, 64), (64, 128, 128), (128, 128, 64), (4096, 128, 128), (1000, 1000, 2048), (100000, 64, 64)]:
    run(f"nt {c}", lambda c=c: nt_case(*c))
for c in [(8, 56, 56, 64, 64, 1, 1), (8, 56, 56, 64, 64, 3, 1), (8, 56, 56, 128, 128, 3, 2), (8, 56, 56, 256, 512, 1, 2),
          (32, 14, 14, 256, 256, 3, 1), (128, 7, 7, 512, 512, 3, 1)]:
    conv_case(*c)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(results, open("gpurun_out/diag_gemm.json", "w"), indent=1)
print("SUMMARY ok=%d bad=%d" % (sum(r["ok"] for r in results), sum(not r["ok"] for r in results)))
