#!/usr/bin/env python
"""Small workloads for compute-sanitizer (memcheck / racecheck / initcheck) on the GPU box:
  python tools/sanitize_case.py stages   - staged calls, 2 lanes, 320x240, ORDERED then FAST, a few frames
  python tools/sanitize_case.py batch    - BatchedPipeline, 64 lanes, 320x240, in order, a few pushes (the bench's launch shapes)
  python tools/sanitize_case.py batch2   - the same on two streams (overlap=2)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import common, harness as H
from rebvio_b200 import capi, synth
from gpu_impl import GpuImpl

what = sys.argv[1] if len(sys.argv) > 1 else "stages"
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 4
name = sys.argv[3] if len(sys.argv) > 3 else "C5:320x240"
if what == "stages":
    spec, frames, R_rel = common.sequence(name, nf)
    cfg = common.cfgs_from_spec(spec)
    for mode in (capi.REDUCE_ORDERED, capi.REDUCE_FAST):
        g = GpuImpl(cfg[0], cfg[1], cfg[2], cfg[3], lanes=2, lane=1, reduce_mode=mode)
        G = H.run_lean(g, frames, R_rel)
        print("mode", mode, "frames", len(G), "matches", G[-1]["nm"], "Vg", G[-1]["Vg"], flush=True)
        g.close()
else:
    from rebvio_b200.pipeline import BatchedPipeline
    import bench
    B = int(os.environ.get("LANES", "64"))
    spec = synth.config(name, n_frames=nf)
    seq = synth.generate(spec)
    cam, det, mapc, core = bench.cfg_structs(spec)
    u8 = np.stack([np.stack([np.ascontiguousarray(seq.frames[(k + b) % nf]) for b in range(B)]) for k in range(nf)]).astype(np.uint8)
    R = np.stack([np.stack([seq.R_rel[k] for b in range(B)]) for k in range(nf)]).astype(np.float32)
    lag = 2 if what == "batch2" else 0
    pipe = BatchedPipeline(cam, det, mapc, core, B, overlap=lag)
    for k in range(nf):
        o = pipe.push(frames=u8[k], **({"R_rel": R[k - lag]} if k >= lag + 1 else {}))
        print(k, {key: np.asarray(v).ravel()[:2].tolist() for key, v in o.items() if key in ("keylines", "matches", "Vg")}, flush=True)
    pipe.ctx.sync()
    pipe.close()
print("done")
