"""First GPU bring-up: teacher-forced and free-running parity of the CUDA path against the oracle (and the compiled
reference when oracle/_ref is loadable). Prints mismatch stats; not a test."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from rebvio_b200 import synth, capi
from oracle import pyref
import harness as H
from gpu_impl import GpuImpl

name = sys.argv[1] if len(sys.argv) > 1 else "C2"
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 5
spec = synth.config(name, n_frames=nf)
seq = synth.generate(spec)
frames = [synth.to_float3(f) for f in seq.frames]
cam, det, mapc, core = pyref.default_cfgs(cam=dict(fm=spec.cam.fm, cx=spec.cam.cx, cy=spec.cam.cy, rows=spec.cam.rows, cols=spec.cam.cols),
                                          det=dict(keylines_ref=spec.keylines_ref, keylines_max=spec.keylines_max))
orc = pyref.OracleLib(cam, det, core, mapc)
t = time.time(); T = H.run_stages(orc, frames, seq.R_rel); print("oracle run", round(time.time() - t, 2), "s")
try:
    ref = pyref.RefLib(cam, det, core)
    Rr = H.run_stages(ref, frames, seq.R_rel, teacher=None)
    print("oracle vs ref (free-run, X differs by SVD backend):", H.dirty(H.report(Rr, T))[:6])
except Exception as e:
    print("reference .so not usable here:", e)
for mode, mname in ((capi.REDUCE_ORDERED, "ORDERED"), (capi.REDUCE_FAST, "FAST")):
    for teacher in (T, None):
        g = GpuImpl(cam, det, mapc, core, lanes=2, lane=1, reduce_mode=mode)
        t = time.time()
        G = H.run_stages(g, frames, seq.R_rel, teacher=teacher)
        rep = H.report(T, G)
        bad = H.dirty(rep)
        print(f"== {mname} {'teacher-forced' if teacher is not None else 'free-running'}: {len(bad)} dirty stages, {round(time.time()-t,2)} s, X_rel max",
              max(f['extrot']['X_rel'] for f in rep[1:]))
        for b in bad[:25]:
            print("   ", b)
        g.close()
