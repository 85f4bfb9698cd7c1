"""Per-shape conv benchmark: this library's tcgen05 implicit GEMM vs cuDNN (torch, channels_last bf16).
Device-timed (CUDA events, L2 flushed between iterations). Writes gpurun_out/conv_bench.json."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from pytorch_distributed_training_tutorials_b200.ops import native as NV
from pytorch_distributed_training_tutorials_b200.utils.timing import time_cuda

NB = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
# (H, Cin, Cout, R, stride, count) : the 22 distinct non-stem convs of ResNet-50 at 224^2
SHAPES = [(56, 64, 64, 1, 1, 1), (56, 64, 64, 3, 1, 3), (56, 64, 256, 1, 1, 4), (56, 256, 64, 1, 1, 2),
          (56, 256, 128, 1, 1, 1), (56, 128, 128, 3, 2, 1), (28, 128, 512, 1, 1, 4), (56, 256, 512, 1, 2, 1),
          (28, 512, 128, 1, 1, 3), (28, 128, 128, 3, 1, 3), (28, 512, 256, 1, 1, 1), (28, 256, 256, 3, 2, 1),
          (14, 256, 1024, 1, 1, 6), (28, 512, 1024, 1, 2, 1), (14, 1024, 256, 1, 1, 5), (14, 256, 256, 3, 1, 5),
          (14, 1024, 512, 1, 1, 1), (14, 512, 512, 3, 2, 1), (7, 512, 2048, 1, 1, 3), (14, 1024, 2048, 1, 2, 1),
          (7, 2048, 512, 1, 1, 2), (7, 512, 512, 3, 1, 2)]
rows = []
tot = {"ours": [0, 0, 0], "cudnn": [0, 0, 0]}
print(f"NB={NB}  | shape | GFLOP | ours fprop/dgrad/wgrad ms (TFLOP/s) | cudnn fprop/dgrad/wgrad ms")
for (H, Cin, Cout, R, s, cnt) in SHAPES:
    x = torch.randn(NB, H, H, Cin, device=dev).bfloat16()
    w = (torch.randn(Cout, R, R, Cin, device=dev) * 0.05).bfloat16()
    Ho = H // s
    dy = torch.randn(NB, Ho, Ho, Cout, device=dev).bfloat16()
    gf = 2.0 * NB * Ho * Ho * Cout * Cin * R * R / 1e9
    t_f = time_cuda(lambda: NV.conv2d_fprop(x, w, stride=s), iters=10)
    t_d = time_cuda(lambda: NV.conv2d_dgrad(dy, w, x.shape, R, stride=s), iters=10)
    t_w = time_cuda(lambda: NV.conv2d_wgrad(x, dy, R, stride=s), iters=10)
    xc = x.permute(0, 3, 1, 2); wc = w.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last); dyc = dy.permute(0, 3, 1, 2)
    c_f = time_cuda(lambda: F.conv2d(xc, wc, stride=s, padding=R // 2), iters=10)
    c_d = time_cuda(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [s, s], [R // 2, R // 2], [1, 1], False, [0, 0], 1, [True, False, False]), iters=10)
    c_w = time_cuda(lambda: torch.ops.aten.convolution_backward(dyc, xc, wc, None, [s, s], [R // 2, R // 2], [1, 1], False, [0, 0], 1, [False, True, False]), iters=10)
    for i, v in enumerate((t_f, t_d, t_w)):
        tot["ours"][i] += v * cnt
    for i, v in enumerate((c_f, c_d, c_w)):
        tot["cudnn"][i] += v * cnt
    rows.append(dict(H=H, Cin=Cin, Cout=Cout, R=R, stride=s, count=cnt, gflop=gf, ours=[t_f, t_d, t_w], cudnn=[c_f, c_d, c_w]))
    print(f"{H:3d} {Cin:4d}->{Cout:4d} R{R} s{s} x{cnt} | {gf:7.1f} | {t_f:.3f} ({gf / t_f:.0f}) {t_d:.3f} ({gf / t_d:.0f}) {t_w:.3f} ({gf / t_w:.0f}) | "
          f"{c_f:.3f} {c_d:.3f} {c_w:.3f}", flush=True)
print("weighted totals (ms/step): ours fprop/dgrad/wgrad = %.2f %.2f %.2f | cudnn = %.2f %.2f %.2f" % (*tot["ours"], *tot["cudnn"]))
# plain GEMM peak check
for (M, N, K) in [(8192, 8192, 8192), (200704, 256, 64), (50176, 512, 512), (12544, 2048, 512)]:
    a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
    t = time_cuda(lambda: NV.gemm_tn(a, b), iters=10)
    c = time_cuda(lambda: a @ b.t(), iters=10)
    print(f"gemm {M}x{N}x{K}: ours {t:.3f} ms ({2.0 * M * N * K / t / 1e9:.0f} TFLOP/s) cublas {c:.3f} ms ({2.0 * M * N * K / c / 1e9:.0f})")
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"NB": NB, "rows": rows, "totals": tot}, open("gpurun_out/conv_bench.json", "w"), indent=1)
