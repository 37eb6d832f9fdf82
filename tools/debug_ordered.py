#!/usr/bin/env python
"""n < keylines_max probe of the ORDERED tryVel / extRotVel paths: addend planes and sums of the chain-kernel form against the first form
(RBV_TUNE_ORDERED = 1) and the oracle, both lanes. usage (GPU box): python tools/debug_ordered.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import common, harness as H
from oracle import pyref
from rebvio_b200 import capi
from gpu_impl import GpuImpl

rows, cols = 240, 320
kmax = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
spec, frames, R_rel = common.sequence("C5:320x240", 3)
cfg = common.cfgs(spec.cam.fm, spec.cam.cx, spec.cam.cy, rows, cols, int(0.75 * kmax), kmax)
orc = pyref.OracleLib(cfg[0], cfg[1], cfg[3], cfg[2])
a, b = orc.detect(frames[0]), orc.detect(frames[1])
orc.build_distance_field(b)
n = orc.map_size(a)
print("kmax", kmax, "n_old", n, "n_new", orc.map_size(b))
res = np.zeros(n, np.float32)
ref = {}
for vel in ([0, 0, 0], [0.01, -0.02, 0.005]):
    ref[tuple(vel)] = orc.try_vel(a, np.float32(vel), 1e3, res.copy())
planes = {}
for name, tune in (("legacy", {capi.TUNE_ORDERED: 1, capi.TUNE_TRYVEL: 1}), ("chain2", {capi.TUNE_TRYVEL: 2}), ("chain4", {capi.TUNE_TRYVEL: 4})):
    g = GpuImpl(cfg[0], cfg[1], cfg[2], cfg[3], lanes=2, lane=1, reduce_mode=capi.REDUCE_ORDERED, tune=tune)
    ga, gb = g.detect(frames[0]), g.detect(frames[1])
    g.build_distance_field(gb)
    g.ctx.resetResiduals(ga)
    for vel in ([0, 0, 0], [0.01, -0.02, 0.005]):
        sc, JtJ, JtF = g.ctx.tryVel(ga, vel, 1e3)
        so, JJo, JFo = ref[tuple(vel)]
        print(name, vel, "score lanes", sc, "oracle", so, "JtJ00 lanes", JtJ[:, 0], "oracle", JJo.ravel()[0], "bit-equal lane1:",
              H.bits(np.float32(sc[1])) == H.bits(np.float32(so)), np.array_equal(H.bits(JtJ[1]), H.bits(JJo)))
        P = np.zeros((2, 10, n), np.float32)
        for lane in range(2):
            for pl in range(10):
                buf = np.zeros(n, np.float32)
                capi.check(g.ctx.L.rbv_debug_read_addends(g.ctx.h, pl, lane, capi.ptr(buf), n))
                P[lane, pl] = buf
        planes[(name, tuple(vel))] = P
        print("   lanes' planes equal:", np.array_equal(H.bits(P[0]), H.bits(P[1])), " plane sums (sequential float32) lane1:",
              [float(np.add.accumulate(P[1, pl], dtype=np.float32)[-1]) for pl in (0, 1)])
    g.close()
for vel in ([0, 0, 0], [0.01, -0.02, 0.005]):
    L, C2, C4 = planes[("legacy", tuple(vel))], planes[("chain2", tuple(vel))], planes[("chain4", tuple(vel))]
    for nm, X in (("chain2", C2), ("chain4", C4)):
        bad = np.argwhere(H.bits(L[1]) != H.bits(X[1]))
        print(vel, nm, "vs legacy planes: mismatching entries", len(bad), bad[:6].tolist())

# ---- per-chunk running sums of accumulator 0 as the chain kernel saw them (RBV_DEBUG_CHAIN=1) against numpy prefix sums
os.environ["RBV_DEBUG_CHAIN"] = "1"
g = GpuImpl(cfg[0], cfg[1], cfg[2], cfg[3], lanes=2, lane=1, reduce_mode=capi.REDUCE_ORDERED, tune={capi.TUNE_TRYVEL: 4})
ga, gb = g.detect(frames[0]), g.detect(frames[1])
g.build_distance_field(gb)
g.ctx.resetResiduals(ga)
sc, JtJ, JtF = g.ctx.tryVel(ga, [0, 0, 0], 1e3)
print("chain4 again: score lanes", sc)
CH = 512
nch = (n + CH - 1) // CH
for lane in range(2):
    buf = np.zeros(n, np.float32); dbg = np.zeros(nch, np.float32)
    capi.check(g.ctx.L.rbv_debug_read_addends(g.ctx.h, 0, lane, capi.ptr(buf), n))
    capi.check(g.ctx.L.rbv_debug_read_addends(g.ctx.h, -1, lane, capi.ptr(dbg), nch))
    pre = np.add.accumulate(buf, dtype=np.float32)
    want = np.array([pre[min((c + 1) * CH, n) - 1] for c in range(nch)], np.float32)
    print("lane", lane, "chunk: kernel running sum / numpy prefix / delta of this chunk (kernel, numpy)")
    prevk = prevw = np.float32(0)
    for c in range(nch):
        flag = "" if H.bits(dbg[c]) == H.bits(want[c]) else "  <-- differs"
        print("  ", c, dbg[c], want[c], dbg[c] - prevk, want[c] - prevw, flag)
        prevk, prevw = dbg[c], want[c]
g.close()
