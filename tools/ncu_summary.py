#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`): per-launch time, DRAM bytes, throughputs, occupancy, top stalls.
usage: python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep > profiles/x_summary.md"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
def g(r, k, d=""):
    return r[col[k]] if k in col else d
stall_keys = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
print(f"# ncu --set full summary of {rep}\n")
print("| kernel | grid | block | time us | dram rd MB | dram wr MB | dram % | L2 MB | sm busy % | issue % | warps active % | regs | top stalls (warps per issue) |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows[2:]:
    name = g(r, "Kernel Name").split("(")[0]
    st = sorted(((float(r[col[k]] or 0), k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for k in stall_keys), reverse=True)[:3]
    def mb(k):
        v, u = float(g(r, k, "0") or 0), units[col[k]] if k in col else ""
        return v * {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6}.get(u, 1.0)
    print(f"| {name} | {g(r,'Grid Size')} | {g(r,'Block Size')} | {float(g(r,'gpu__time_duration.sum','0')):.1f} | {mb('dram__bytes_read.sum'):.1f} | {mb('dram__bytes_write.sum'):.1f} | "
          f"{float(g(r,'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','0')):.1f} | {mb('lts__t_bytes.sum'):.0f} | {float(g(r,'sm__throughput.avg.pct_of_peak_sustained_elapsed','0')):.1f} | "
          f"{float(g(r,'smsp__issue_active.avg.pct_of_peak_sustained_active','0')):.1f} | {float(g(r,'sm__warps_active.avg.pct_of_peak_sustained_active','0')):.1f} | {g(r,'launch__registers_per_thread')} | "
          + ", ".join(f"{n} {v:.2f}" for v, n in st) + " |")
