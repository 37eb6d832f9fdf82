"""Shared helpers of the parity tests: golden fixtures, config construction, implementation factories."""
from __future__ import annotations

import os

import numpy as np

import golden_io
from oracle import pyref
from rebvio_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["g160", "g131", "g320"]

# Tolerance on extRotVel's X (6-vector): the reference solves the 6x6 system with LAPACK sgesvd_ in float32
# (TooN::SVD<6,6,float>, core.cpp:247-248; LAPACK is an un-vendored dependency), the oracle and the CUDA path with a
# double-precision Jacobi eigen-solve. A backward-stable float32 solve is only accurate to eps32 * cond(Wx), so that is
# the bound (floor 1e-5 = the north star's pose tolerance); measured: 3e-6 .. 2e-4 for cond 44 .. 7900, always
# below eps32 * cond. Everything else on the path is compared bit for bit.
X_RTOL_FLOOR = 1e-5
# CUDA vs oracle: same algorithm, same statement order, FP64 -> agreement to float32 rounding of the result
X_RTOL_VS_ORACLE = 1e-5


def load_golden(name):
    frames_u8, R_rel, meta, recs, extra = golden_io.load(os.path.join(GOLDEN, name + ".npz"))
    frames = [synth.to_float3(f) for f in frames_u8]
    return frames, R_rel, meta, recs, extra


def cfgs(fm, cx, cy, rows, cols, kref, kmax, **core_over):
    return pyref.default_cfgs(cam=dict(fm=fm, cx=cx, cy=cy, rows=int(rows), cols=int(cols)),
                              det=dict(keylines_ref=int(kref), keylines_max=int(kmax)), core=core_over or None)


def cfgs_from_meta(meta):
    fm, cx, cy, rows, cols = meta["cam"]
    return cfgs(fm, cx, cy, rows, cols, meta["kref"], meta["kmax"])


def cfgs_from_spec(spec):
    return cfgs(spec.cam.fm, spec.cam.cx, spec.cam.cy, spec.cam.rows, spec.cam.cols, spec.keylines_ref, spec.keylines_max)


_SEQ_CACHE = {}


def sequence(name, n_frames):
    """cached synthetic sequence -> (spec, float frames, R_rel)"""
    key = (name, n_frames)
    if key not in _SEQ_CACHE:
        spec = synth.config(name, n_frames=n_frames)
        seq = synth.generate(spec)
        _SEQ_CACHE[key] = (spec, [synth.to_float3(f) for f in seq.frames], seq.R_rel)
    return _SEQ_CACHE[key]


def x_rel_max(rep):
    return max((f["extrot"]["X_rel"] for f in rep[1:]), default=0.0)


def x_tol_vs_ref(Wx):
    """tolerance on ||X - X_ref|| / ||X_ref|| for the reference's float32 LAPACK solve of Wx X = JtF"""
    cond = np.linalg.cond(np.asarray(Wx, np.float64))
    return max(X_RTOL_FLOOR, float(np.finfo(np.float32).eps) * cond)


def assert_x_vs_ref(ref_recs, rep):
    for k in range(1, len(rep)):
        tol = x_tol_vs_ref(ref_recs[k]["extrot"]["Wx"])
        assert rep[k]["extrot"]["X_rel"] <= tol, (k, rep[k]["extrot"]["X_rel"], tol)


def sequence_long(name, n_frames):
    """long sequences (hundreds of frames) for the GPU tests: the same scene generator, rendered by synth.render_batch_torch on the
    device when there is one (the numpy renderer takes ~0.3 s per frame) -> (spec, float frames, R_rel). Both sides of a comparison
    are fed these same frames, so the +-1 grey level between the two renderers does not matter."""
    key = ("long", name, n_frames)
    if key not in _SEQ_CACHE:
        import torch
        spec = synth.config(name, n_frames=n_frames)
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        frames_u8, R_rel = synth.batched_sequences(spec, [spec.seed], n_frames, device=dev, n_textures=1)
        f = frames_u8[:, 0].cpu().numpy()
        _SEQ_CACHE[key] = (spec, [synth.to_float3(f[k]) for k in range(n_frames)], R_rel[:, 0])
    return _SEQ_CACHE[key]
