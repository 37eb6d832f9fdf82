#!/usr/bin/env python
"""Per-kernel instruction / stall hot spots of an .ncu-rep (source page). usage: ncu_hot.py rep kernel [min_instr_k] [launch]"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# several launches are concatenated: split at header rows
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Address":
        hdr = r; cur = []; blocks.append(cur)
    elif cur is not None and r and r[0].startswith("0x") and len(r) >= len(hdr):
        cur.append(r)
data = blocks[int(sys.argv[4]) if len(sys.argv) > 4 else 0]
ci = {h: i for i, h in enumerate(hdr)}
I, S, T = ci["Instructions Executed"], ci["# Samples"], ci["Thread Instructions Executed"]
ti, tt, ts = sum(int(r[I]) for r in data), sum(int(r[T]) for r in data), sum(int(r[S]) for r in data)
print(f"{kern}: {len(blocks)} launches; warp instr {ti/1e6:.1f}M thread instr {tt/1e6:.1f}M (avg active {tt/max(ti,1):.1f}) samples {ts} sass {len(data)}")
ops = collections.Counter()
for r in data:
    w = r[ci["Source"]].split()
    op = w[1] if w[0].startswith("@") else w[0]
    ops[op.split(".")[0]] += int(r[I])
print("opcode mix (M warp instr):", ", ".join(f"{o} {n/1e6:.1f}" for o, n in ops.most_common(18)))
if thr:
    for k, r in enumerate(data):
        if int(r[I]) >= thr * 1000 or int(r[S]) * 200 > ts:
            print(k, r[I].rjust(9), r[T].rjust(11), r[S].rjust(5), r[ci["Source"]].strip()[:90])
