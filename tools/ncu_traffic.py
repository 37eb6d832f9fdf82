#!/usr/bin/env python
"""DRAM traffic per launch (dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured launches of each kernel) from an
`ncu --set full` report -> JSON that bench.py reads for `roofline.traffic`.
usage: python tools/ncu_traffic.py gpurun_out/prof_x.ncu-rep "<captured command>" > profiles/rNN_ncu_traffic.json"""
import csv, io, json, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
acc = {}
for r in rows[2:]:
    name = r[col["Kernel Name"]].split("(")[0]
    b = sum(float(r[col[k]] or 0) * scale.get(units[col[k]], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    a = acc.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += b
print(json.dumps({"source": sys.argv[1], "command": sys.argv[2] if len(sys.argv) > 2 else "",
                  "note": "mean dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full --clock-control none",
                  "bytes_per_launch": {n: t / c for n, (c, t) in sorted(acc.items())},
                  "launches_captured": {n: c for n, (c, t) in sorted(acc.items())}}, indent=1))
