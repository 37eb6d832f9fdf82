#!/usr/bin/env python
"""Cycles per keyline of the ORDERED chain kernels at a known keyline count (GPU box): python tools/chain_timing.py [kmax] [lanes]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import common
from rebvio_b200 import capi
from gpu_impl import GpuImpl

rows, cols = 240, 320
kmax = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
lanes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
spec, frames, R_rel = common.sequence("C5:320x240", 3)
cfg = common.cfgs(spec.cam.fm, spec.cam.cx, spec.cam.cy, rows, cols, int(0.75 * kmax), kmax)
for mode, tune in (("one launch", {capi.TUNE_TRYVEL: 4}), ("two kernels", {capi.TUNE_TRYVEL: 4, capi.TUNE_ORDERED: 2})):
    g = GpuImpl(cfg[0], cfg[1], cfg[2], cfg[3], lanes=lanes, lane=0, reduce_mode=capi.REDUCE_ORDERED, tune=tune)
    ga, gb = g.detect(frames[0]), g.detect(frames[1])
    n = g.map_size(ga)
    g.build_distance_field(gb)
    g.ctx.resetResiduals(ga)
    tot = 0.0
    for name in ("k_tryvel_ord4", "k_chain_tryvel", "k_extrot_ord", "k_chain_extrot"):
        g.ctx.profile_set(name); g.ctx.profile_reset()
        for _ in range(30):
            if "tryvel" in name: g.ctx.tryVel(ga, [0.01, -0.02, 0.005], 1e3)
            else:
                g.ctx.forwardMatch(ga, gb); g.ctx.extRotVel(gb, np.zeros(3, np.float32))
        r = g.ctx.profile_read().get(name)
        if not r or not r[1]: continue
        us = 1e3 * r[0] / r[1]
        print(f"[{mode}] {name}: lanes {lanes} n {n}: {us:.2f} us per launch = {us * 1965.0 / max(n, 1):.2f} cycles per keyline at 1965 MHz")
    g.ctx.profile_set(-1)
    g.close()
