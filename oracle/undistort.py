"""CPU restatement of cv::undistort as Rebvio::imageCallback uses it (TEST INFRASTRUCTURE, like the rest of oracle/).

Reference call sites: rebvio/src/rebvio.cpp:43-45 (convertTo(CV_32F, 3.0) then Camera::undistort), rebvio/include/rebvio/camera.hpp:26-40,54-58
(K = [fm 0 cx; 0 fm cy; 0 0 1] and D = (k1, k2, p1, p2, k3), both float32). The arithmetic lives in OpenCV, an un-vendored dependency
("OpenCV3", README.md:16, no version pinned); this file restates the published algorithm of cv::undistort (OpenCV 4.x
calib3d/undistort.dispatch.cpp: per stripe initUndistortRectifyMap -> CV_16SC2 fixed-point maps with INTER_BITS = 5 -> remap INTER_LINEAR,
BORDER_CONSTANT 0) and is pinned against the Python build of the same library (cv2 4.13.0 in the build container) by
tests/golden/undistort_*.npz (tools/make_undistort_golden.py). See DESIGN.md 10 for the parity statement."""
from __future__ import annotations

import numpy as np

INTER_BITS = 5
INTER_TAB_SIZE = 1 << INTER_BITS


def fixed_point_map(rows, cols, fx, fy, cx, cy, dist):
    """-> (sx, sy int32 [rows, cols], wx, wy int32 0..31): source pixel and 1/32-pixel fraction of every destination pixel"""
    fx, fy, cx, cy = (float(np.float32(v)) for v in (fx, fy, cx, cy))          # camera.hpp:39: K_ is a float32 matrix
    k1, k2, p1, p2, k3 = (float(np.float32(v)) for v in dist)                    # camera.hpp:40: D_ likewise
    j = np.arange(cols, dtype=np.float64)[None, :]
    i = np.arange(rows, dtype=np.float64)[:, None]
    x = j * (1.0 / fx) + (-cx * (1.0 / fx))                                      # iR = inv(Ar): [1/fx 0 -cx/fx; 0 1/fy -cy/fy; 0 0 1]
    y = i * (1.0 / fy) + (-cy * (1.0 / fy))
    x = np.broadcast_to(x, (rows, cols)); y = np.broadcast_to(y, (rows, cols))
    x2, y2 = x * x, y * y
    r2 = x2 + y2
    _2xy = 2 * x * y
    kr = 1 + ((k3 * r2 + k2) * r2 + k1) * r2
    xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2)
    yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy
    u = fx * xd + cx
    v = fy * yd + cy
    iu = np.rint(u * INTER_TAB_SIZE).astype(np.int64)                            # saturate_cast<int> = cvRound (half to even)
    iv = np.rint(v * INTER_TAB_SIZE).astype(np.int64)
    return ((iu >> INTER_BITS).astype(np.int32), (iv >> INTER_BITS).astype(np.int32),
            (iu & (INTER_TAB_SIZE - 1)).astype(np.int32), (iv & (INTER_TAB_SIZE - 1)).astype(np.int32))


def undistort(img_f32, fx, fy, cx, cy, dist):
    """img_f32 [rows, cols] float32 (0..765) -> undistorted float32, cv::remap bilinear with the float weight table"""
    img = np.ascontiguousarray(img_f32, np.float32)
    rows, cols = img.shape
    sx, sy, wx, wy = fixed_point_map(rows, cols, fx, fy, cx, cy, dist)
    scale = np.float32(1.0 / INTER_TAB_SIZE)
    t = np.arange(INTER_TAB_SIZE, dtype=np.float32) * scale                      # interpolateLinear: (1 - x, x) in float32
    c0, c1 = (np.float32(1.0) - t), t
    w00 = c0[wy] * c0[wx]; w01 = c0[wy] * c1[wx]; w10 = c1[wy] * c0[wx]; w11 = c1[wy] * c1[wx]   # tab = vy[k1] * vx[k2]

    def fetch(yy, xx):
        ok = (yy >= 0) & (yy < rows) & (xx >= 0) & (xx < cols)
        out = np.zeros((rows, cols), np.float32)                                 # BORDER_CONSTANT, value 0
        out[ok] = img[yy[ok], xx[ok]]
        return out
    s00, s01, s10, s11 = fetch(sy, sx), fetch(sy, sx + 1), fetch(sy + 1, sx), fetch(sy + 1, sx + 1)
    return ((s00 * w00 + s01 * w01) + s10 * w10) + s11 * w11                     # float32, left to right, separate roundings
