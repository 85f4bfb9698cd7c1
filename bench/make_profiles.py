"""Assemble the committed evidence under profiles/ from the raw files that gpurun merged into gpurun_out/."""
import csv, collections, glob, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
HBM, TF = peaks["hbm_gbs"], peaks["bf16_tflops"]


def launches():
    f = os.path.join(G, "launches.csv")
    if not os.path.exists(f):
        return
    lines = [l for l in open(f) if not l.startswith("==")]
    tot = collections.defaultdict(float); cnt = collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", "")); ns = v * {"ns": 1, "us": 1e3, "ms": 1e6}.get(row["Metric Unit"], 1)
        k = re.sub(r"\(.*", "", row["Kernel Name"])[:80]
        tot[k] += ns; cnt[k] += 1
    total = sum(tot.values())
    out = ["# Kernel launch list of ONE ResNet-50 training step (batch 256, 224x224, 1 GPU, eager launches)", "",
           "`ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off` around exactly one step of",
           "`bench/profile_step.py` (after the same command exited 0 without ncu). Serialised, cold-cache durations:",
           "compare shares, not absolutes.", "", f"total kernel time {total / 1e6:.2f} ms over {sum(cnt.values())} launches", "",
           "| kernel | launches | ms | share |", "|---|---|---|---|"]
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:20]:
        own = "b200::" in k or k.startswith("zero_phase")
        out.append(f"| `{k}`{'' if own else ' (torch)'} | {cnt[k]} | {v / 1e6:.3f} | {100 * v / total:.1f}% |")
    open(os.path.join(P, "launches_resnet50_step.md"), "w").write("\n".join(out) + "\n")


def conv_bench():
    f = os.path.join(G, "conv_bench.json")
    if not os.path.exists(f):
        return
    d = json.load(open(f))
    out = [f"# ResNet-50 convolutions, batch {d['NB']}: tcgen05 implicit GEMM (ours) vs cuDNN (torch channels_last bf16)", "",
           "CUDA events, L2 flushed between iterations, median of 10 (`bench/conv_bench.py`). TFLOP/s = 2*MACs / time;",
           f"`of measured` = fraction of the measured cuBLAS bf16 burst peak ({TF:.0f} TFLOP/s, MEASURED_PEAKS.json).", "",
           "| H | Cin->Cout | R | s | x | GFLOP | ours fprop ms (TFLOP/s, of measured) | ours dgrad | ours wgrad | cuDNN fprop / dgrad / wgrad ms |", "|---|---|---|---|---|---|---|---|---|---|"]
    for r in d["rows"]:
        o, c, g = r["ours"], r["cudnn"], r["gflop"]
        cell = lambda t: f"{t:.3f} ({g / t:.0f}, {g / t / TF * 100:.0f}%)"
        out.append(f"| {r['H']} | {r['Cin']}->{r['Cout']} | {r['R']} | {r['stride']} | {r['count']} | {g:.1f} | {cell(o[0])} | {cell(o[1])} | {cell(o[2])} | {c[0]:.3f} / {c[1]:.3f} / {c[2]:.3f} |")
    t = d["totals"]
    out += ["", f"Weighted per-step totals (ms): ours fprop/dgrad/wgrad = {t['ours'][0]:.2f} / {t['ours'][1]:.2f} / {t['ours'][2]:.2f}; "
            f"cuDNN = {t['cudnn'][0]:.2f} / {t['cudnn'][1]:.2f} / {t['cudnn'][2]:.2f}."]
    open(os.path.join(P, "conv_bench_vs_cudnn.md"), "w").write("\n".join(out) + "\n")


def sweeps():
    out = ["# All-reduce sweep: fused peer-memory kernels (ours) vs NCCL 2.28.9", "",
           "`bench/allreduce_sweep.py`, CUDA events, max over ranks. Ours = ONE kernel doing gather -> scale -> (cast) -> reduce over",
           "NVLink peer memory -> scatter, on an ordinary (non-symmetric) tensor like NCCL's input. busbw = algbw * 2(N-1)/N.",
           "`inplace` = the tensor lives in the symmetric arena (what DDP's gradient views use; like NCCL user-buffer registration):",
           "no copy-in / copy-out, the reduce phase alone.",
           "NVLink roofline: 770 GB/s measured peer copy per direction per GPU (B200_PROFILING.md); the multicast (NVLS) path",
           "could not be exercised: torch's symmetric memory reports no multicast support on these boxes.", ""]
    inputs = sorted(glob.glob(os.path.join(G, "allreduce_sweep_*.json")))
    if not inputs:
        return                                   # keep the committed sweep: nothing newer to build it from
    for f in inputs:
        d = json.load(open(f)); n = d["world"]; fac = 2 * (n - 1) / n
        keys = [k for k in ("nccl_ms", "oneshot_ms", "twoshot_ms", "nvls_ms", "twoshot_inplace_ms", "nvls_inplace_ms",
                            "twoshot_bf16wire_ms", "nvls_bf16wire_ms") if any(k in r for r in d["rows"])]
        out += [f"## {n} GPUs (multicast available: {d['multicast']})", "", "| bytes | " + " | ".join(k[:-3] + " us / busbw GB/s" for k in keys) + " |", "|---|" + "---|" * len(keys)]
        for r in d["rows"]:
            out.append(f"| {r['bytes']} | " + " | ".join((f"{r[k] * 1e3:.1f} / {r['bytes'] / (r[k] * 1e-3) * fac / 1e9:.1f}" if k in r else "-") for k in keys) + " |")
        out.append("")
    open(os.path.join(P, "allreduce_sweep.md"), "w").write("\n".join(out) + "\n")


def benches():
    out = ["# bench.py results (ResNet-50 DDP bf16, synthetic 224x224, per-GPU batch 256)", "",
           "| file | impl | GPUs | img/s (device-timed) | ms/step | e2e img/s | SM MHz (median) | throttle reasons | kernel launches in timed region |", "|---|---|---|---|---|---|---|---|---|"]
    # rows of earlier runs whose raw files are gone (every gpurun box is fresh) are kept; a new file of the same
    # name replaces its row
    rows = {}
    prev = os.path.join(P, "bench_results.md")
    if os.path.exists(prev):
        for l in open(prev):
            cells = [c.strip() for c in l.strip().strip("|").split("|")]
            if len(cells) == 9 and cells[0].endswith(".json"):
                rows[cells[0]] = l.rstrip("\n")
    for f in sorted(glob.glob(os.path.join(G, "scale_*.json")) + glob.glob(os.path.join(G, "bench_*.json"))):
        try:
            line = [l for l in open(f) if l.startswith("{")][-1]; d = json.loads(line)
        except Exception:
            continue
        if "value" not in d:
            continue
        c = d.get("clocks") or {}
        rows[os.path.basename(f)] = (f"| {os.path.basename(f)} | {d.get('impl')} | {d['n_gpus']} | {d['value']:.0f} | {d['ms_per_step']:.2f} | {d['e2e']['value']:.0f} | {c.get('sm_mhz')} | {c.get('reasons')} | {d.get('gpu_launches')} |")
    out += [rows[k] for k in sorted(rows)]
    open(prev, "w").write("\n".join(out) + "\n")


def misc():
    for name in ("graphed_pipeline_bench.json", "pipeline_bench.json", "mp_bench.json", "toy_bench_dp_1.json", "toy_bench_dp_2.json", "toy_bench_ddp_2.json", "toy_bench_dp_4.json", "toy_bench_dp_8.json", "toy_bench_ddp_8.json", "conv_bench.json", "diag_gemm.json"):
        f = os.path.join(G, name)
        if os.path.exists(f):
            json.dump(json.load(open(f)), open(os.path.join(P, name), "w"), indent=1)
    for name in ("mp_bench.log",):
        f = os.path.join(G, name)
        if os.path.exists(f):
            open(os.path.join(P, name.replace(".log", ".md")), "w").write("".join(l for l in open(f) if l.startswith("|")))
    for name in ("ddp_overhead_2.txt", "ddp_overhead_8.txt"):
        f = os.path.join(G, name)
        if os.path.exists(f):
            open(os.path.join(P, name), "w").write(open(f).read())
    f = os.path.join(G, "graphed_pipeline_bench.json")
    if os.path.exists(f):
        d = json.load(open(f))
        out = [f"# ResNet-50 split over 2 GPUs, mini-batch {d['batch']} at {d['size']}x{d['size']}: naive vs pipelined (BASELINE config 4)", "",
               "`bench/graphed_pipeline_bench.py`: zero_grad + forward + loss + backward + SGD per step, device-timed on both GPUs.",
               "`graphed` = every (stage, micro-batch) replayed from CUDA graphs (`parallel/graphed_pipeline.py`).", "",
               "| variant | ms/step | img/s |", "|---|---|---|"]
        for k, v in d.items():
            if k.endswith("_chosen"):
                out.append(f"| {k[:-7]}: time-balanced cut chosen after `{v}` | | |")
            if k.endswith("_ms") and isinstance(v, (int, float)):
                out.append(f"| {k[:-3]} | {v:.2f} | {d['batch'] / v * 1e3:.0f} |")
        open(os.path.join(P, "pipeline_bench.md"), "w").write("\n".join(out) + "\n")
    for b in (8, 64):
        f = os.path.join(G, f"launches_{b}.csv")
        if os.path.exists(f):
            txt = subprocess.run([sys.executable, os.path.join(ROOT, "bench", "agg_launches.py"), f], capture_output=True, text=True).stdout
            open(os.path.join(P, f"launches_resnet50_step_bs{b}.md"), "w").write(
                f"# Launch list of one ResNet-50 step at batch {b} (where the per-launch fixed costs show)\n\n"
                f"`ncu --metrics gpu__time_duration.sum` around one eager step of `bench/profile_step.py {b}`. Compare with the\n"
                "batch-256 list: a kernel whose time does not fall with the batch is latency-bound.\n\n```\n" + txt + "```\n")


def ncu_reports():
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max"]
    out = ["# ncu --set full captures (raw metrics of interest)", "", "Read with `ncu -i <rep> --page raw --csv`; one row per captured launch.", ""]
    for rep in sorted(glob.glob(os.path.join(G, "*.ncu-rep"))):
        try:
            txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, timeout=120).stdout
        except Exception:
            continue
        rows = list(csv.reader(txt.splitlines()))
        if len(rows) < 3:
            continue
        hdr = rows[0]
        out += [f"## {os.path.basename(rep)}", "", "| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(rows) - 2)) + " |", "|---|---|" + "---|" * (len(rows) - 2)]
        for i, h in enumerate(hdr):
            if h in keys or h == "Kernel Name":
                out.append(f"| {h} | {rows[1][i]} | " + " | ".join(r[i][:60] for r in rows[2:]) + " |")
        out.append("")
    open(os.path.join(P, "ncu_summaries.md"), "w").write("\n".join(out) + "\n")


launches(); conv_bench(); sweeps(); benches(); misc()      # ncu: bench/ncu_summarise.py
print(sorted(os.listdir(P)))
