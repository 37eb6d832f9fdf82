"""Camera::undistort: the oracle restatement of cv::undistort against the committed golden vectors of the
build container's OpenCV (CPU), and the CUDA remap against the oracle and the same golden vectors (GPU). Bit-exact."""
import os
import zlib

import numpy as np
import pytest

import common
from oracle import undistort as U
from rebvio_b200 import capi, host, synth

GOLD = os.path.join(common.GOLDEN, "undistort_euroc.npz")


def _inputs():
    rng = np.random.RandomState(77)
    noise = rng.randint(0, 256, (480, 752)).astype(np.uint8)
    scene = np.ascontiguousarray(synth.generate(synth.config("C2", n_frames=1)).frames[0]).astype(np.uint8)
    return {"noise": noise, "scene": scene}


def _check_against_golden(name, img, g):
    crc = np.array([zlib.crc32(np.ascontiguousarray(img[r]).tobytes()) for r in range(img.shape[0])], np.uint32)
    assert np.array_equal(crc, g[name + "_crc"]), f"{name}: rows differ from cv2.undistort: {np.nonzero(crc != g[name + '_crc'])[0][:8]}"
    assert np.array_equal(img[g["rows"]].view(np.uint32), g[name + "_rows"].view(np.uint32))


def test_oracle_undistort_matches_opencv_golden():
    g = np.load(GOLD)
    K4, D5 = g["K4"], g["D5"]
    for name, u8 in _inputs().items():
        assert zlib.crc32(u8.tobytes()) == int(g[name + "_u8_crc"][0]), "synthetic input generator changed: regenerate the golden"
        out = U.undistort(u8.astype(np.float32) * np.float32(3.0), K4[0], K4[1], K4[2], K4[3], D5)
        _check_against_golden(name, out, g)


def test_oracle_undistort_identity_and_border():
    """zero distortion is the identity; a strongly distorted model reads outside the image -> BORDER_CONSTANT 0"""
    rng = np.random.RandomState(1)
    img = rng.randint(0, 256, (60, 80)).astype(np.float32) * np.float32(3.0)
    same = U.undistort(img, 70.0, 70.0, 40.0, 30.0, np.zeros(5, np.float32))
    assert np.array_equal(same, img)
    far = U.undistort(img, 70.0, 70.0, 40.0, 30.0, np.array([0.9, 0.5, 0.0, 0.0, 0.0], np.float32))
    assert (far[0, :5] == 0).all() and (far[30, 38:42] != 0).any()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["u8", "f32", "u8_prefetch"])
def test_cuda_undistort_matches_oracle_and_golden(kind):
    g = np.load(GOLD)
    K4, D5 = g["K4"], g["D5"]
    ins = _inputs()
    cam, det, mapc, core = capi.defaults()
    lanes = 2
    ctx = host.Context(cam, det, mapc, core, lanes=lanes)
    ctx.set_undistort(K4, D5)
    frames = np.stack([ins["noise"], ins["scene"]])
    if kind == "f32":
        src = frames.astype(np.float32) * np.float32(3.0)
        m, n, at = ctx.detect(src, np.full(lanes, det.threshold, np.float32))
    else:
        ptrs = (capi.C.c_void_p * lanes)(*[frames.ctypes.data + b * frames.strides[0] for b in range(lanes)])
        if kind == "u8_prefetch":
            capi.check(ctx.L.rbv_prefetch_frames(ctx.h, ptrs, 0, 1))
        m = ctx.new_map()
        n = np.zeros(lanes, np.int32); at = np.zeros(lanes, np.float32)
        ts = np.zeros(lanes, np.uint64); thr = np.full(lanes, det.threshold, np.float32)
        capi.check(ctx.L.rbv_detect_u8(ctx.h, m.h, ptrs, 0, capi.ptr(ts), capi.ptr(thr), capi.ptr(n), capi.ptr(at)))
    for b, name in enumerate(("noise", "scene")):
        got = ctx.read_image(4, lane=b)
        exp = U.undistort(ins[name].astype(np.float32) * np.float32(3.0), K4[0], K4[1], K4[2], K4[3], D5)
        assert np.array_equal(got.view(np.uint32), exp.view(np.uint32)), name
        _check_against_golden(name, got, g)
    assert n[1] > 1000                                   # the front end ran on the undistorted scene
    assert ctx.profile_launches().get("k_undistort", 0) == 1 and ctx.profile_launches().get("k_u8_to_f32", 0) == 0
    # switching it off restores the plain conversion path
    ctx.set_undistort(None)
    m2, n2, _ = ctx.detect(frames.astype(np.float32) * np.float32(3.0), np.full(lanes, det.threshold, np.float32))
    assert np.array_equal(ctx.read_image(4, lane=0), frames[0].astype(np.float32) * np.float32(3.0))
    ctx.close()
