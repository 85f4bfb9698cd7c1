"""Per-shape bandwidth of the BN kernels (CUDA events, L2 flushed)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_distributed_training_tutorials_b200.ops import native as NV
from pytorch_distributed_training_tutorials_b200.utils.timing import time_cuda
dev = torch.device("cuda:0")
NB = 256
for (H, C) in [(112, 64), (56, 64), (56, 256), (28, 128), (28, 512), (14, 256), (14, 1024), (7, 512), (7, 2048)]:
    rows = NB * H * H
    x = torch.randn(rows, C, device=dev).bfloat16(); dy = torch.randn(rows, C, device=dev).bfloat16(); res = torch.randn(rows, C, device=dev).bfloat16()
    g = torch.ones(C, device=dev); b = torch.zeros(C, device=dev); rm = torch.zeros(C, device=dev); rv = torch.ones(C, device=dev)
    mb = rows * C * 2 / 1e6
    t_s = time_cuda(lambda: NV.bn_stats(x), iters=10)
    sums = NV.bn_stats(x)
    t_a = time_cuda(lambda: NV.bn_apply(x, None, sums, g, b, rm, rv, 1e-5, 0.1, True), iters=10)
    y, mean, invstd = NV.bn_apply(x, None, sums, g, b, rm, rv, 1e-5, 0.1, True)
    t_b = time_cuda(lambda: NV.bn_backward(dy, None, x, mean, invstd, g, True, False, beta=b), iters=10)
    t_br = time_cuda(lambda: NV.bn_backward(dy, y, x, mean, invstd, g, True, True, beta=b), iters=10)
    print(f"{H:3d}x{H:<3d} C={C:4d} {mb:6.0f} MB | stats {t_s*1e3:6.0f} us ({mb/t_s/1e3:.2f} TB/s) | apply {t_a*1e3:6.0f} us ({2*mb/t_a/1e3:.2f} TB/s) | "
          f"bwd {t_b*1e3:6.0f} us ({7*mb/t_b/1e3:.2f} TB/s over 7 passes) | bwd+res {t_br*1e3:6.0f} us ({7*mb/t_br/1e3:.2f} TB/s)")
