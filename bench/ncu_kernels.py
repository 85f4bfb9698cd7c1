"""One kernel family per invocation, a handful of launches: the target of `ncu --set full` (bench/ncu_all.sh).

    python bench/ncu_kernels.py <family>

Prints the CUDA-event time of the measured launches so the plain run (which must precede every ncu run) already
gives the un-profiled duration that the roofline fractions in profiles/ncu_summaries.md are computed from."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pytorch_distributed_training_tutorials_b200 import _lib
from pytorch_distributed_training_tutorials_b200.ops import native as NV

dev = torch.device("cuda:0")
NB = int(os.environ.get("NCU_NB", "256"))      # batch; NCU_NB=8 exposes the latency floor of a launch
bf = torch.bfloat16


def conv_inputs(H, Cin, Cout, R, s):
    x = torch.randn(NB, H, H, Cin, device=dev).to(bf)
    w = (torch.randn(Cout, R, R, Cin, device=dev) * 0.05).to(bf)
    dy = torch.randn(NB, H // s, H // s, Cout, device=dev).to(bf)
    return x, w, dy


def family(name):
    """-> (callable, algorithmic bytes per launch, FLOPs per launch)"""
    if name.startswith("conv_"):
        _, what, H, Cin, Cout, R, s = name.split("_")
        H, Cin, Cout, R, s = map(int, (H, Cin, Cout, R, s))
        x, w, dy = conv_inputs(H, Cin, Cout, R, s)
        flops = 2.0 * NB * (H // s) ** 2 * Cout * Cin * R * R
        bx, bw_, by = x.numel() * 2, w.numel() * 2, dy.numel() * 2
        if what == "fprop":
            sums = NV.bn_sums_buffer(Cout, dev)
            return (lambda: NV.conv2d_fprop(x, w, stride=s, bn_sums=sums)), bx + bw_ + by, flops
        if what == "dgrad":
            return (lambda: NV.conv2d_dgrad(dy, w, x.shape, R, stride=s)), by + bw_ + bx, flops
        if what == "dgradacc":
            acc = torch.randn_like(x)
            return (lambda: NV.conv2d_dgrad(dy, w, x.shape, R, stride=s, accumulate_into=acc)), by + bw_ + 2 * bx, flops
        if what == "wgrad":
            return (lambda: NV.conv2d_wgrad(x, dy, R, stride=s)), bx + by + w.numel() * 4, flops
    if name == "linear_8192x4096x4096":
        a = torch.randn(8192, 4096, device=dev).to(bf); b = torch.randn(4096, 4096, device=dev).to(bf)
        return (lambda: NV.gemm_tn(a, b)), (a.numel() + b.numel() + 8192 * 4096) * 2, 2.0 * 8192 * 4096 * 4096
    if name.startswith("bn_"):
        _, what, H, C = name.split("_")
        H, C = int(H), int(C)
        rows = NB * H * H
        x = torch.randn(rows, C, device=dev).to(bf); res = torch.randn(rows, C, device=dev).to(bf)
        dy = torch.randn(rows, C, device=dev).to(bf)
        g = torch.rand(C, device=dev) + 0.5; b = torch.randn(C, device=dev) * 0.1
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        sums = NV.bn_stats(x)
        mask = torch.empty(rows * C // 8, device=dev, dtype=torch.uint8)
        e = rows * C
        if what == "applyres":      # bn3: + residual, ReLU, bit mask
            return (lambda: NV.bn_apply(x, res, sums, g, b, rm, rv, 1e-5, 0.1, True, None, mask)), e * 6 + e // 8, 0
        if what == "apply":         # bn1 / bn2: ReLU, no residual
            return (lambda: NV.bn_apply(x, None, sums, g, b, rm, rv, 1e-5, 0.1, True)), e * 4, 0
        y, mean, invstd = NV.bn_apply(x, res, sums, g, b, rm, rv, 1e-5, 0.1, True, None, mask)
        if what == "bwdres":        # reduce<3> + apply<0> (masked gradient materialised for the skip branch)
            return (lambda: NV.bn_backward(dy, None, x, mean, invstd, g, True, True, beta=b, relu_mask=mask)), e * 12 + e // 8, 0
        if what == "bwd":           # reduce<2> + apply<2> (mask recomputed from x)
            return (lambda: NV.bn_backward(dy, None, x, mean, invstd, g, True, False, beta=b)), e * 10, 0
        if what == "stats":
            return (lambda: NV.bn_stats(x)), e * 2, 0
    if name == "maxpool":
        x = torch.randn(NB, 112, 112, 64, device=dev).to(bf)
        y, arg = NV.maxpool_fwd(x)
        dyp = torch.randn_like(y)
        return (lambda: (NV.maxpool_fwd(x), NV.maxpool_bwd(dyp, arg, x.shape))), x.numel() * 2 + y.numel() * 3 + y.numel() * 3 + x.numel() * 2, 0
    if name == "avgpool":
        x = torch.randn(NB, 7, 7, 2048, device=dev).to(bf)
        dyp = torch.randn(NB, 2048, device=dev).to(bf)
        return (lambda: (NV.avgpool_fwd(x), NV.avgpool_bwd(dyp, x.shape))), x.numel() * 4, 0
    if name == "stem":
        x = torch.randn(NB, 3, 224, 224, device=dev)
        return (lambda: NV.stem_im2col(x)), x.numel() * 4 + NB * 112 * 112 * 160 * 2, 0
    if name in ("stem_implicit", "stem_implicit_fwd", "stem_implicit_wgrad"):
        x = torch.randn(NB, 3, 224, 224, device=dev); w = torch.randn(64, 7, 7, 3, device=dev) * 0.1
        sums = NV.bn_sums_buffer(64, dev)
        dy = torch.randn(NB, 112, 112, 64, device=dev).to(bf)
        fl = 2.0 * NB * 112 * 112 * 64 * 147
        if name == "stem_implicit_fwd":
            return (lambda: NV.stem_conv_fprop(x, w, bn_sums=sums)), x.numel() * 4 + dy.numel() * 2, fl
        if name == "stem_implicit_wgrad":
            return (lambda: NV.stem_conv_wgrad(x, dy)), x.numel() * 4 + dy.numel() * 2, fl
        return (lambda: (NV.stem_conv_fprop(x, w, bn_sums=sums), NV.stem_conv_wgrad(x, dy))), \
            2 * x.numel() * 4 + 2 * dy.numel() * 2, 2 * fl
    if name == "stem_explicit":
        x = torch.randn(NB, 3, 224, 224, device=dev)
        wk = torch.randn(64, 160, device=dev).to(bf)
        sums = NV.bn_sums_buffer(64, dev)
        dy = torch.randn(NB * 112 * 112, 64, device=dev).to(bf)

        def run():
            cols = NV.stem_im2col(x)
            NV.gemm_tn(cols, wk, bn_sums=sums)
            NV.gemm_nt_reduce(dy, cols)
        return run, 3 * NB * 112 * 112 * 160 * 2 + x.numel() * 4 + 2 * dy.numel() * 2, 2 * 2.0 * NB * 112 * 112 * 64 * 160
    if name == "sgd":
        from pytorch_distributed_training_tutorials_b200.ops.optim import FusedSGD
        ps = [torch.randn(2048, 1024, device=dev, requires_grad=True) for _ in range(12)]
        for p in ps:
            p.grad = torch.randn_like(p)
        opt = FusedSGD(ps, lr=1e-3)
        n = sum(p.numel() for p in ps)
        return (lambda: opt.step()), n * (4 + 4 + 4 + 2), 0
    if name == "loss_ce":
        logits = torch.randn(NB, 1000, device=dev); lab = torch.randint(0, 1000, (NB,), device=dev)
        return (lambda: NV.ce_index_fwd_bwd(logits, lab)), logits.numel() * 8, 0
    if name == "allreduce_local":
        # the collective kernel with world == 1 (ncu profiles one process on one GPU): gather -> stage -> scatter
        import ctypes
        from pytorch_distributed_training_tutorials_b200.ops import collectives as coll
        n = 16 << 20
        x = torch.randn(n, device=dev)
        stage = torch.zeros(n * 4 + 4096, dtype=torch.uint8, device=dev)
        sig = torch.zeros(coll.SIGNAL_WORDS, dtype=torch.int32, device=dev)
        table = coll.TensorTable([n], dev, False, False, target_blocks=128)
        table.bind([x], torch.cuda.current_stream(dev))
        status = torch.zeros(4, dtype=torch.int32, device=dev)
        bufs = _lib.u64_array([stage.data_ptr()]); sigs = _lib.u64_array([sig.data_ptr()])

        def run():
            _lib.check(_lib.lib().b200dt_coll_launch(1, bufs, sigs, 0, 0, 1, 0, table.tensors.data_ptr(), table.chunks.data_ptr(),
                                                     table.n_chunks, 0, 0, 0, 1.0, 0, table.grid(128), status.data_ptr(), 0,
                                                     torch.cuda.current_stream(dev).cuda_stream), "coll")
        return run, n * 4 * 4, 0
    raise SystemExit(f"unknown family {name}")


FAMILIES = ["conv_fprop_56_64_256_1_1", "conv_fprop_56_64_64_3_1", "conv_fprop_14_256_256_3_1", "conv_fprop_28_128_128_3_2",
            "conv_dgrad_56_64_64_3_1", "conv_dgradacc_56_256_64_1_1", "conv_dgrad_56_128_128_3_2", "conv_wgrad_28_128_128_3_1",
            "conv_wgrad_56_64_256_1_1", "linear_8192x4096x4096", "bn_applyres_56_256", "bn_apply_56_64", "bn_bwdres_56_256",
            "bn_bwd_56_64", "bn_stats_28_512", "maxpool", "avgpool", "stem", "stem_implicit_fwd", "stem_implicit_wgrad", "stem_explicit", "sgd", "loss_ce", "allreduce_local"]

if __name__ == "__main__":
    name = sys.argv[1]
    if name == "list":
        print(" ".join(FAMILIES)); raise SystemExit
    fn, nbytes, flops = family(name)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[1]
    # the region ncu captures (--profile-from-start off): exactly one more call, cold L2
    flush.zero_()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    fn()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print(json.dumps({"family": name, "ms": ms, "bytes": nbytes, "flops": flops,
                      "GBps": nbytes / ms / 1e6, "TFLOPs": flops / ms / 1e9}), flush=True)
