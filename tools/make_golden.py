"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/librebvio_ref.so, compiled from
/root/reference by `make -C oracle ref`). Run in the build container only; the fixtures travel, the reference does not.

    python tools/make_golden.py

Each fixture holds the u8 input frames, the relative rotations fed to the harness, the camera / detector config and
the per-stage records of tests/harness.run_stages (free-running reference), flattened to "<frame>/<stage>/<...>" keys.
Cases:
  g160   160x120, 3 frames, default-ish config                      (plain case)
  g131   131x97 (ragged: not a multiple of 32 / of the tile sizes), keylines_max = 300 so the keylines_max cut of
         edge_detector.cpp:113-119 is hit, 3 frames
  g320   320x240, 2 frames, C5 configuration of SURVEY.md 8(d)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

from rebvio_b200 import synth
from oracle import pyref
import harness as H
import golden_io

CASES = {
    "g160": dict(cols=160, rows=120, n_frames=3, seed=11, kmax=2000, kref=1500),
    "g131": dict(cols=131, rows=97, n_frames=3, seed=12, kmax=300, kref=250),
    "g320": dict(cols=320, rows=240, n_frames=2, seed=5, kmax=-(-320 * 240 // 22), kref=int(0.75 * -(-320 * 240 // 22))),
}


def main():
    for name, c in CASES.items():
        spec = synth.SequenceSpec(name, synth.CameraSpec(0.6 * c["cols"], c["cols"] / 2.0, c["rows"] / 2.0, c["rows"], c["cols"]),
                                  c["n_frames"], c["seed"], keylines_ref=c["kref"], keylines_max=c["kmax"])
        seq = synth.generate(spec)
        frames = [synth.to_float3(f) for f in seq.frames]
        cam, det, mapc, core = pyref.default_cfgs(
            cam=dict(fm=spec.cam.fm, cx=spec.cam.cx, cy=spec.cam.cy, rows=spec.cam.rows, cols=spec.cam.cols),
            det=dict(keylines_ref=spec.keylines_ref, keylines_max=spec.keylines_max))
        ref = pyref.RefLib(cam, det, core)
        recs = H.run_stages(ref, frames, seq.R_rel)
        s0, s1, dog, mag = ref.scale_space(frames[0])
        ref.close()
        meta = dict(cam=[cam.fm, cam.cx, cam.cy, cam.rows, cam.cols], kref=det.keylines_ref, kmax=det.keylines_max)
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        golden_io.save(path, np.stack(seq.frames), np.asarray(seq.R_rel, np.float64), meta, recs,
                       extra={} if name == "g320" else {"scale/s0": s0, "scale/s1": s1, "scale/dog": dog, "scale/mag": mag})
        print(name, "keylines", [r["detect"]["n"] for r in recs], "matches", [r.get("directed", {}).get("n") for r in recs],
              os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
