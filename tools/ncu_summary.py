"""Summarise .ncu-rep files into a Markdown table (duration, DRAM bytes / throughput, tensor-pipe activity, registers)
against the measured peaks in MEASURED_PEAKS.json.  python tools/ncu_summary.py gpurun_out/prof_*.ncu-rep > profiles/ncu_summary.md"""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
src = "measured" if (ROOT / "MEASURED_PEAKS.json").exists() else "fallback"
KEYS = {
    "dur_us": "gpu__time_duration.sum",
    "dram_rd": "dram__bytes_read.sum",
    "dram_wr": "dram__bytes_write.sum",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "tensor_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "issue_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread",
    "smem": "launch__shared_mem_per_block_dynamic",
    "grid": "launch__grid_size",
    "warps_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
}
print(f"# ncu summaries (peaks: HBM {peaks['hbm_gbs']} GB/s, bf16 {peaks['bf16_tflops']} TFLOP/s burst, of {src})\n")
print("Captured with `ncu --set full --clock-control none --import-source on` under gpurun (one GPU), after the same command exited 0 without ncu.")
print("Durations are under the profiler (serialised, cold caches): compare shares, not absolutes; the timed numbers live in the bring-up logs.\n")
print("| report | kernel | grid | regs | dyn smem | duration us | DRAM read MB | DRAM write MB | DRAM GB/s (frac of peak) | DRAM % | tensor pipe % | issue active % | warps active % |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        def get(k):
            name = KEYS[k]
            if name in hdr:
                i = hdr.index(name)
                try:
                    v = float(r[i].replace(",", ""))
                except ValueError:
                    return None
                u = units[i]
                if k == "dur_us":
                    v = v / 1e3 if u in ("nsecond", "ns") else (v * 1e3 if u in ("msecond", "ms") else v)
                if k in ("dram_rd", "dram_wr"):
                    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                    v = v * mult / 1e6
                return v
            return None
        d = {k: get(k) for k in KEYS}
        name = r[hdr.index("Kernel Name")]
        name = name.replace("void b200::", "").split("(")[0][:70]
        bw = ((d["dram_rd"] or 0) + (d["dram_wr"] or 0)) * 1e6 / (d["dur_us"] * 1e-6) / 1e9 if d["dur_us"] else 0
        f = lambda v, n=1: "-" if v is None else f"{v:.{n}f}"
        print(f"| {Path(rep).name} | `{name}` | {f(d['grid'],0)} | {f(d['regs'],0)} | {f(d['smem'],0)} | {f(d['dur_us'])} | {f(d['dram_rd'])} | "
              f"{f(d['dram_wr'])} | {bw:.0f} ({bw / peaks['hbm_gbs']:.2f}) | {f(d['dram_pct'])} | {f(d['tensor_pct'])} | {f(d['issue_pct'])} | {f(d['warps_pct'])} |")
