"""Read gpurun_out/ncu/*.ncu-rep (captured by bench/ncu_all.sh) with `ncu -i ... --page raw --csv` and write
profiles/ncu_summaries.md: one row per captured kernel launch with the roofline fractions."""
import csv
import glob
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
HBM, TF = peaks["hbm_gbs"], peaks["bf16_tflops"]
plain = {}
for line in open(os.path.join(ROOT, "gpurun_out/ncu/plain.jsonl")):
    try:
        d = json.loads(line)
        plain[d["family"]] = d
    except Exception:
        pass

# families captured in an earlier call of the round (every gpurun box is fresh and ncu_all.sh starts a new plain.jsonl):
# their un-profiled time is carried over from the table this script wrote last time
prev_md = os.path.join(ROOT, "profiles/ncu_summaries.md")
if os.path.exists(prev_md):
    for line in open(prev_md):
        cells = [c.strip() for c in line.strip().strip("|").split("|")]
        if len(cells) == 12 and cells[0] not in plain and cells[0] not in ("family", "---"):
            try:
                plain[cells[0]] = {"family": cells[0], "ms": float(cells[4]), "flops": 0}
            except ValueError:
                pass

WANT = {"gpu__time_duration.sum": "dur", "dram__bytes_read.sum": "rd", "dram__bytes_write.sum": "wr",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pct",
        "sm__inst_executed_pipe_tensor.sum": "tensor_inst",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_pct",
        "launch__registers_per_thread": "regs", "launch__grid_size": "grid", "launch__block_size": "block",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
        "lts__t_sector_hit_rate.pct": "l2_hit",
        "smsp__cycles_active.avg": "cyc"}


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return None


rows = []
for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out/ncu/*.ncu-rep"))):
    fam = os.path.basename(rep)[:-8]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    if len(rd) < 3:
        continue
    hdr, units = rd[0], rd[1]
    for r in rd[2:]:
        rec = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        row = {"family": fam, "kernel": rec.get("Kernel Name", "?")}
        for k, short in WANT.items():
            if k in rec:
                v = num(rec[k])
                if v is not None and short in ("dur",):
                    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u[k], 1e-6)      # -> ms
                if v is not None and short in ("rd", "wr"):
                    v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u[k], 1)
                row[short] = v
        rows.append(row)

out = ["# ncu --set full, one capture per kernel family (current binaries)", "",
       "`bench/ncu_all.sh` on one B200: every family of `bench/ncu_kernels.py` is run plainly first (CUDA-event time below,",
       "cold L2, batch 256 shapes of the ResNet-50 step), then once more under",
       "`ncu --set full --clock-control none --import-source on --profile-from-start off` with the measured call bracketed by",
       "cudaProfilerStart/Stop. `of HBM` = DRAM bytes ncu counted / ncu duration / measured copy bandwidth "
       f"({HBM:.0f} GB/s, MEASURED_PEAKS.json); `of tensor peak` = algorithmic FLOPs / un-profiled duration / measured cuBLAS bf16 burst ({TF:.0f} TFLOP/s).",
       "Durations under ncu are serialised and a little longer than the plain ones; fractions use the column they name.",
       "The raw `.ncu-rep` files are not committed (60 MB); `bench/ncu_all.sh` regenerates them.", "",
       "| family | kernel | grid x block | regs | plain ms (family) | ncu ms | DRAM rd+wr MB | of HBM (measured) | tensor pipe active % | of tensor peak (measured) | warps active % | L2 hit % |",
       "|---|---|---|---|---|---|---|---|---|---|---|---|"]
for r in rows:
    p = plain.get(r["family"], {})
    dur = r.get("dur") or 0
    traffic = (r.get("rd") or 0) + (r.get("wr") or 0)
    hbm_frac = traffic / (dur * 1e-3) / (HBM * 1e9) if dur else 0
    tf_frac = (p.get("flops", 0) / (p.get("ms", 1) * 1e-3) / (TF * 1e12)) if p.get("flops") else None
    k = r["kernel"]
    k = k[:60]
    out.append(f"| {r['family']} | `{k}` | {int(r.get('grid') or 0)} x {int(r.get('block') or 0)} | {int(r.get('regs') or 0)} | "
               f"{p.get('ms', float('nan')):.3f} | {dur:.3f} | {traffic / 1e6:.0f} | {100 * hbm_frac:.0f}% | "
               f"{(r.get('tensor_pct') or 0):.1f} | {('%.0f%%' % (100 * tf_frac)) if tf_frac is not None else '-'} | "
               f"{(r.get('warps_pct') or 0):.0f} | {(r.get('l2_hit') or 0):.0f} |")
open(os.path.join(ROOT, "profiles/ncu_summaries.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
