"""Run one conv shape a few times (for ncu --set full on tc_gemm_kernel).
usage: one_conv.py NB H Cin Cout R stride [fprop|dgrad|dgrad_acc|wgrad]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_distributed_training_tutorials_b200.ops import native as NV
NB, H, Cin, Cout, R, s = map(int, sys.argv[1:7])
what = sys.argv[7] if len(sys.argv) > 7 else "fprop"
dev = torch.device("cuda:0")
x = torch.randn(NB, H, H, Cin, device=dev).bfloat16()
w = (torch.randn(Cout, R, R, Cin, device=dev) * 0.05).bfloat16()
dy = torch.randn(NB, H // s, H // s, Cout, device=dev).bfloat16()
acc = torch.randn(NB, H, H, Cin, device=dev).bfloat16()
fn = {"fprop": lambda: NV.conv2d_fprop(x, w, stride=s), "dgrad": lambda: NV.conv2d_dgrad(dy, w, x.shape, R, stride=s),
      "dgrad_acc": lambda: NV.conv2d_dgrad(dy, w, x.shape, R, stride=s, accumulate_into=acc),
      "wgrad": lambda: NV.conv2d_wgrad(x, dy, R, stride=s)}[what]
for _ in range(5):
    fn()
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    fn()
e1.record(); torch.cuda.synchronize()
print(f"{what} {sys.argv[1:7]}: {e0.elapsed_time(e1) / 10:.4f} ms")
