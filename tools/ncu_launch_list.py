#!/usr/bin/env python
"""Summarise the ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X.csv <cmd>`).
usage: python tools/ncu_launch_list.py gpurun_out/launches_x.csv "<command>" > profiles/x_ncu_launch_list.md"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
cmd = sys.argv[2] if len(sys.argv) > 2 else ""
agg = {}
for r in rows:
    name = r[4].split("(")[0]
    ns = float(r[14]) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[13], 1.0)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ns
tot = sum(a[1] for a in agg.values())
print(f"# ncu launch list (gpu__time_duration.sum, --clock-control none) of `{cmd}`\n")
print(f"Only this library's kernels (-k regex:^k_); {len(rows)} launches, {tot/1e6:.2f} ms total. Per-launch times are cold-cache and "
      f"serialised: compare SHARES with the CUDA-event table, not absolutes.\n")
print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {n} | {c} | {t/1e3:.1f} | {t/1e3/c:.1f} | {t/tot:.3f} |")
