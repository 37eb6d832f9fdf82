#!/usr/bin/env python
"""ptxas -v resources and SASS mnemonic counts of every kernel of the built objects (runs on the CPU box: cuobjdump only).
usage: python tools/sass_summary.py > profiles/rNN_sass_summary.md"""
import glob, os, re, subprocess, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "rebvio_b200", "csrc", "build")
res = {}
cur = None
for line in (l for f in sorted(glob.glob(os.path.join(BUILD, "*.ptxas.log"))) for l in open(f)):
    m = re.search(r"Compiling entry function '(\w+)'", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.split("(")[0].strip(); res[cur] = {}
    m = re.search(r"Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", line)
    if m and cur: res[cur].update(regs=int(m.group(1)), bars=int(m.group(2) or 0), stack=int(m.group(3) or 0), smem=int(m.group(4) or 0))
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m and cur: res[cur].update(spill=int(m.group(2)) + int(m.group(3)))
WATCH = ("UBLKCP", "SYNCS", "LDGSTS", "LDG", "STG", "LDS", "STS", "ATOMS", "ATOMG", "RED", "REDUX", "SHFL", "BAR", "FADD", "FMUL", "FFMA", "DADD", "DFMA", "MUFU", "NANOSLEEP", "ELECT")
ops = {}
for obj in sorted(glob.glob(os.path.join(BUILD, "*.o"))):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    k = None
    for line in out.splitlines():
        m = re.search(r"Function : (\w+)", line)
        if m:
            k = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.split("(")[0].strip(); ops[k] = collections.Counter(); continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and k: ops[k][m.group(1)] += 1; ops[k]["_total"] += 1
print("# SASS / ptxas summary of the built kernels (sm_100a, nvcc 12.9, -fmad=false -lineinfo)\n")
print("`python tools/sass_summary.py`: registers, barriers, static shared memory, spills from `ptxas -v`; instruction counts from `cuobjdump -sass`.")
print("UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier operations, LDGSTS = cp.async, FFMA / DFMA appear only inside the IEEE division, reciprocal and square-root sequences (-fmad=false: no a*b+c contraction anywhere).\n")
print("| kernel | regs | static smem B | barriers | spill B | SASS instr | " + " | ".join(WATCH) + " |")
print("|---|---|---|---|---|---|" + "---|" * len(WATCH))
for k in sorted(k for k in ops if k):
    r = res.get(k, {})
    print(f"| {k} | {r.get('regs','')} | {r.get('smem','')} | {r.get('bars','')} | {r.get('spill','')} | {ops[k]['_total']} | " + " | ".join(str(ops[k].get(w, 0) or "") for w in WATCH) + " |")
