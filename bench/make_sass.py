"""Write the committed SASS evidence: docs/sass/<kernel>.sass for the hot kernels (instruction text only,
encodings stripped) and docs/sass/MNEMONICS.md with the counts of the Blackwell-specific opcodes."""
import os, re, subprocess, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "pytorch_distributed_training_tutorials_b200", "_native", "libb200dt.so")
out_dir = os.path.join(ROOT, "docs", "sass"); os.makedirs(out_dir, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = {}
cur = None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); funcs[cur] = []; continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\*", line)
    if m:
        funcs[cur].append(f"/*{m.group(1)}*/ {m.group(2).strip()} ;")
keep = {"tc_gemm_kernel": "tc_gemm_kernel", "coll_kernelILi1ELi8": "coll_twoshot_world8", "coll_kernelILi2": "coll_nvls",
        "coll_kernelILi0": "coll_oneshot", "coll_kernelILi3": "coll_broadcast", "bn_bwd_reduce_kernelILi2": "bn_bwd_reduce_remask",
        "bn_apply_kernelILb1": "bn_apply_train", "multi_opt_kernelILb0": "multi_sgd", "p2p_bcast_kernel": "p2p_broadcast",
        "p2p_reduce_kernel": "p2p_reduce", "mse_static_kernelIf": "mse_static",
        "skinny_fwd_kernel": "skinny_linear_fwd", "relu_bwd_bias_kernel": "relu_bwd_bias",
        "stem_im2col_kernel": "stem_im2col", "maxpool_bwd_kernel": "maxpool_bwd", "maxpool_fwd_kernel": "maxpool_fwd",
        "stem_fprop_kernel": "stem_fprop_implicit", "stem_wgrad_kernel": "stem_wgrad_implicit",
        "p2p_multi_copy_kernel": "p2p_multi_copy", "bn_bwd_apply_kernelILi0": "bn_bwd_apply",
        "multi_opt_kernelILb1": "multi_adam", "splitk_fold_kernel": "splitk_fold", "ce_kernelIfLi1": "cross_entropy"}
pats = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "SYNCS", "LDGMC", "STGMC", "ATOMG", "REDG", "HMMA", "ELECT", "UBLKCP", "ACQBULK", "PREEXIT", "BRA.U.ANY"]
rows = []
for name, body in funcs.items():
    for key, short in keep.items():
        if key in name:
            open(os.path.join(out_dir, short + ".sass"), "w").write(f"// {name}\n" + "\n".join(body) + "\n")
    cnt = collections.Counter()
    for ins in body:
        for p in pats:
            if re.search(r"\b" + re.escape(p), ins):
                cnt[p] += 1
    if cnt:
        rows.append((name, len(body), cnt))
md = ["# SASS mnemonics per kernel (cuobjdump -sass of _native/libb200dt.so, sm_100a)", "",
      "`UTCHMMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `UTCATOMSWS` = tcgen05.alloc/dealloc,",
      "`UTMALDG` / `UTMASTG` / `UTMAREDG` = TMA load / store / reduce-add, `SYNCS` = mbarrier, `LDGMC` / `STGMC`-class =",
      "multimem (NVLS), `ATOMG...SYS` = system-scope flag CAS, `ACQBULK` / `PREEXIT` = griddepcontrol.wait / .launch_dependents",
      "(programmatic dependent launch), `BRA.U.ANY` = back-edge of a waterfall loop around a uniform-datapath instruction",
      "(tc_gemm_kernel had one per UTCHMMA / UTMALDG before its issuing warps ran converged with an elected lane).",
      "No `HMMA` (legacy mma.sync) anywhere.", "",
      "| kernel | instructions | " + " | ".join(pats) + " |", "|---|---|" + "---|" * len(pats)]
for name, n, cnt in sorted(rows, key=lambda r: -r[1]):
    md.append(f"| `{name[:70]}` | {n} | " + " | ".join(str(cnt.get(p, 0) or "") for p in pats) + " |")
open(os.path.join(out_dir, "MNEMONICS.md"), "w").write("\n".join(md) + "\n")
print(len(funcs), "functions;", sorted(os.listdir(out_dir)))
