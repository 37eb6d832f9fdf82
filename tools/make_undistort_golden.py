#!/usr/bin/env python
"""Golden vectors for Camera::undistort: runs the build container's OpenCV (cv2.undistort, the Python build of the
library the reference links, rebvio/include/rebvio/camera.hpp:54-58) on seeded frames with the reference's hard-coded EuRoC cam0
model (camera.hpp:26-40) and stores, per case, the CRC32 of every output row (whole-image bit-exactness in 2 KB) plus the full values
of a few rows (top / middle / bottom, where the border rule acts). Inputs are regenerated from the seed by the tests.
usage: python tools/make_undistort_golden.py   (needs cv2; writes tests/golden/undistort_euroc.npz)"""
import os, sys, zlib
import numpy as np
import cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rebvio_b200 import synth

FX, FY, CX, CY = 458.654, 457.296, 367.215, 248.375
FM = 0.5 * (FX + FY)
D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0], np.float32)
K = np.array([[FM, 0, CX], [0, FM, CY], [0, 0, 1]], np.float32)


def frames():
    rng = np.random.RandomState(77)
    yield "noise", rng.randint(0, 256, (480, 752)).astype(np.uint8)
    spec = synth.config("C2", n_frames=1)
    yield "scene", np.ascontiguousarray(synth.generate(spec).frames[0]).astype(np.uint8)


out = {"K4": np.array([FM, FM, CX, CY], np.float32), "D5": D, "cv2_version": np.array(cv2.__version__)}
ROWS = [0, 1, 2, 239, 240, 477, 478, 479]
for name, u8 in frames():
    f = u8.astype(np.float32) * np.float32(3.0)                      # rebvio.cpp:43
    ref = cv2.undistort(f, K, D.reshape(1, 5))                       # rebvio.cpp:45 -> camera.hpp:56
    out[name + "_crc"] = np.array([zlib.crc32(np.ascontiguousarray(ref[r]).tobytes()) for r in range(ref.shape[0])], np.uint32)
    out[name + "_rows"] = ref[ROWS].copy()
    out[name + "_u8_crc"] = np.array([zlib.crc32(u8.tobytes())], np.uint32)
out["rows"] = np.array(ROWS, np.int32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "undistort_euroc.npz"), **out)
print("written", {k: v.shape for k, v in out.items()})
