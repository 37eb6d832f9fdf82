"""Host-side mirror of the reference's hot-path class API over the C ABI (Python flavour, used by tests and bench.py;
the C++ flavour with the reference's exact signatures lives in rebvio_b200/cpp/rebvio_b200/).

  Context  ~ rebvio::EdgeDetector + rebvio::Core for B lanes (edge_detector.hpp:39-93, core.hpp:99-205)
  Map      ~ B rebvio::EdgeMap (edge_map.hpp:28-135)
Method names follow the reference members: detect, buildDistanceField, rotateKeylines, minimizeVel, forwardMatch,
extRotVel, directedMatch, regularize1Iter, updateInverseDepth, estimateQuantile.
All per-lane arguments are arrays with leading dimension B.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import KEYLINE_DTYPE, check, ptr


def _f32(a, shape):
    a = np.ascontiguousarray(a, np.float32)
    return a.reshape(shape)


class Map:
    def __init__(self, ctx: "Context"):
        self.ctx = ctx
        h = C.c_void_p()
        check(ctx.L.rbv_map_create(ctx.h, C.byref(h)))
        self.h = h

    def release(self):
        if self.h:
            self.ctx.L.rbv_map_release(self.ctx.h, self.h)
            self.h = None

    def size(self):
        n = np.zeros(self.ctx.lanes, np.int32)
        check(self.ctx.L.rbv_map_size(self.ctx.h, self.h, ptr(n)))
        return n

    def threshold(self):
        t = np.zeros(self.ctx.lanes, np.float32)
        check(self.ctx.L.rbv_map_threshold(self.ctx.h, self.h, ptr(t)))
        return t

    def set_threshold(self, t):
        t = _f32(np.broadcast_to(np.asarray(t, np.float32), (self.ctx.lanes,)), (self.ctx.lanes,))
        check(self.ctx.L.rbv_map_set_threshold(self.ctx.h, self.h, ptr(t)))

    def keylines(self, lane=0):
        """AoS host mirror (EdgeMap::keylines()) + detection pixel per keyline (key of EdgeMap::mask())."""
        cap = self.ctx.det.keylines_max
        kl = np.zeros(cap, KEYLINE_DTYPE)
        px = np.zeros(cap, np.uint32)
        n = C.c_int(0)
        check(self.ctx.L.rbv_map_download(self.ctx.h, self.h, lane, ptr(kl), ptr(px), cap, C.byref(n)))
        return kl[:n.value].copy(), px[:n.value].copy()

    def upload(self, lane, kl, pixel_idx=None):
        kl = np.ascontiguousarray(kl, KEYLINE_DTYPE)
        px = None if pixel_idx is None else np.ascontiguousarray(pixel_idx, np.uint32)
        check(self.ctx.L.rbv_map_upload(self.ctx.h, self.h, lane, ptr(kl), ptr(px), len(kl)))

    def mask(self, lane=0):
        out = np.empty((self.ctx.cam.rows, self.ctx.cam.cols), np.int32)
        check(self.ctx.L.rbv_debug_read_mask(self.ctx.h, self.h, lane, ptr(out)))
        return out

    def residuals(self, lane=0):
        out = np.zeros(self.ctx.det.keylines_max, np.float32)
        check(self.ctx.L.rbv_debug_read_residuals(self.ctx.h, self.h, lane, ptr(out), len(out)))
        return out


class Context:
    def __init__(self, cam=None, det=None, mapc=None, core=None, lanes=1, device=0, reduce_mode=capi.REDUCE_ORDERED):
        self.L = capi.load()
        dc, dd, dm, dk = capi.defaults()
        self.cam, self.det, self.mapc, self.core = cam or dc, det or dd, mapc or dm, core or dk
        self.lanes = lanes
        h = C.c_void_p()
        check(self.L.rbv_create(C.byref(self.cam), C.byref(self.det), C.byref(self.mapc), C.byref(self.core), device, lanes,
                                C.byref(h)))
        self.h = h
        check(self.L.rbv_set_reduce_mode(self.h, reduce_mode))
        # implementation overrides for test / profiling runs (0 = automatic): RBV_TUNE_FIELD, RBV_TUNE_BOX in the environment
        import os
        for key, name in ((capi.TUNE_FIELD, "RBV_TUNE_FIELD"), (capi.TUNE_BOX, "RBV_TUNE_BOX")):
            if os.environ.get(name):
                self.set_tuning(key, int(os.environ[name]))
        for kv in os.environ.get("RBV_TUNE_ENV", "").split(","):   # "key=value,..." (rbv_set_tuning keys), e.g. "8=1"
            if kv:
                self.set_tuning(int(kv.split("=")[0]), int(kv.split("=")[1]))

    def close(self):
        if self.h:
            self.L.rbv_destroy(self.h)
            self.h = None

    def set_reduce_mode(self, mode): check(self.L.rbv_set_reduce_mode(self.h, mode))
    def set_tuning(self, key, value): check(self.L.rbv_set_tuning(self.h, int(key), int(value)))
    def set_undistort(self, K4=None, D5=None):
        """Camera::undistort on the device: K4 = (fx, fy, cx, cy), D5 = (k1, k2, p1, p2, k3); None switches it off"""
        k = None if K4 is None else np.ascontiguousarray(K4, np.float32)
        d = None if D5 is None else np.ascontiguousarray(D5, np.float32)
        check(self.L.rbv_set_undistort(self.h, ptr(k), ptr(d)))

    def set_stream(self, cuda_stream): check(self.L.rbv_set_stream(self.h, C.c_void_p(cuda_stream)))
    def sync(self): check(self.L.rbv_sync(self.h))
    def launch_count(self): return int(self.L.rbv_launch_count(self.h))
    def new_map(self): return Map(self)

    # per-kernel accounting (rbv_profile_*)
    def kernel_names(self):
        return [self.L.rbv_kernel_name(i).decode() for i in range(self.L.rbv_kernel_count())]

    def profile_set(self, mode):
        """mode: -1 off, -2 all kernels, a kernel name or id = that kernel only"""
        if isinstance(mode, str):
            mode = self.kernel_names().index(mode)
        check(self.L.rbv_profile_set(self.h, int(mode)))

    def profile_reset(self): check(self.L.rbv_profile_reset(self.h))

    def profile_read(self):
        """-> {kernel: (total_ms of bracketed launches, bracketed launches, launches since reset)}"""
        out = {}
        for i, name in enumerate(self.kernel_names()):
            ms, nt, nl = C.c_double(0), C.c_ulonglong(0), C.c_ulonglong(0)
            check(self.L.rbv_profile_read(self.h, i, C.byref(ms), C.byref(nt), C.byref(nl)))
            if nl.value:
                out[name] = (ms.value, nt.value, nl.value)
        return out

    def profile_launches(self):
        """-> {kernel: launches since the last profile_reset (or creation)}"""
        return {n: v[2] for n, v in self.profile_read().items()}

    def servo_threshold(self, threshold, last_count):
        return float(self.L.rbv_servo_threshold(C.byref(self.det), float(threshold), int(last_count)))

    # EdgeDetector::detect (edge_detector.cpp:30-43); thresholds already servoed
    def detect(self, images, thresholds, ts_us=None, out_map=None, want_counts=True):
        B = self.lanes
        images = np.ascontiguousarray(images, np.float32).reshape(B, self.cam.rows, self.cam.cols)
        m = out_map or Map(self)
        ptrs = (C.c_void_p * B)(*[images[b].ctypes.data for b in range(B)])
        thr = _f32(np.broadcast_to(np.asarray(thresholds, np.float32), (B,)), (B,))
        ts = np.ascontiguousarray(np.broadcast_to(np.asarray(0 if ts_us is None else ts_us, np.uint64), (B,)))
        n = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
        check(self.L.rbv_detect(self.h, m.h, ptrs, 0, ptr(ts), ptr(thr), ptr(n) if want_counts else None,
                                ptr(at) if want_counts else None))
        return m, n, at

    def detect_device(self, d_images_ptr, thresholds, out_map=None, want_counts=True):
        B = self.lanes
        m = out_map or Map(self)
        thr = _f32(np.broadcast_to(np.asarray(thresholds, np.float32), (B,)), (B,))
        n = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
        check(self.L.rbv_detect_device(self.h, m.h, C.c_void_p(d_images_ptr), None, ptr(thr),
                                       ptr(n) if want_counts else None, ptr(at) if want_counts else None))
        return m, n, at

    def buildDistanceField(self, m): check(self.L.rbv_build_distance_field(self.h, m.h))

    def rotateKeylines(self, m, R):
        R = _f32(np.broadcast_to(np.asarray(R, np.float32).reshape(-1, 9), (self.lanes, 9)), (self.lanes, 9))
        check(self.L.rbv_rotate_keylines(self.h, m.h, ptr(R)))

    def estimateQuantile(self, m, pct=0.9, bins=100):
        out = np.zeros(self.lanes, np.float32)
        check(self.L.rbv_estimate_quantile(self.h, m.h, pct, bins, ptr(out)))
        return out

    def resetResiduals(self, m): check(self.L.rbv_reset_residuals(self.h, m.h))

    def tryVel(self, m, vel, sigma_rho_min):
        B = self.lanes
        v = _f32(np.broadcast_to(np.asarray(vel, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        s = _f32(np.broadcast_to(np.asarray(sigma_rho_min, np.float32), (B,)), (B,))
        JtJ = np.zeros((B, 3, 3), np.float32); JtF = np.zeros((B, 3), np.float32); sc = np.zeros(B, np.float32)
        check(self.L.rbv_try_vel(self.h, m.h, ptr(v), ptr(s), ptr(JtJ), ptr(JtF), ptr(sc)))
        return sc, JtJ, JtF

    def minimizeVel(self, m, vel):
        B = self.lanes
        v = _f32(np.broadcast_to(np.asarray(vel, np.float32).reshape(-1, 3), (B, 3)), (B, 3)).copy()
        Rv = np.zeros((B, 3, 3), np.float32); sc = np.zeros(B, np.float32)
        check(self.L.rbv_minimize_vel(self.h, m.h, ptr(v), ptr(Rv), ptr(sc)))
        return sc, v, Rv

    def lm_trace(self, lane=0):
        out = np.zeros((int(self.core.iterations) + 1, 16), np.float32)
        check(self.L.rbv_debug_read_lm_trace(self.h, lane, ptr(out), out.size))
        return out

    def forwardMatch(self, old, new):
        n = np.zeros(self.lanes, np.int32)
        check(self.L.rbv_forward_match(self.h, old.h, new.h, ptr(n)))
        return n

    def extRotVel(self, new, vel):
        B = self.lanes
        v = _f32(np.broadcast_to(np.asarray(vel, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        Wx = np.zeros((B, 6, 6), np.float32); X = np.zeros((B, 6), np.float32); ok = np.zeros(B, np.int32)
        check(self.L.rbv_ext_rot_vel(self.h, new.h, ptr(v), ptr(Wx), ptr(X), ptr(ok)))
        return ok, Wx, X

    def directedMatch(self, new, old, vel, Rvel, Rback, max_radius=40.0):
        B = self.lanes
        v = _f32(np.broadcast_to(np.asarray(vel, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        Rv = _f32(np.broadcast_to(np.asarray(Rvel, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        Rb = _f32(np.broadcast_to(np.asarray(Rback, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        kf = np.zeros(B, np.int32); n = np.zeros(B, np.int32)
        check(self.L.rbv_directed_match(self.h, new.h, old.h, ptr(v), ptr(Rv), ptr(Rb), max_radius, ptr(kf), ptr(n)))
        return n, kf

    def regularize1Iter(self, m):
        n = np.zeros(self.lanes, np.int32)
        check(self.L.rbv_regularize(self.h, m.h, ptr(n)))
        return n

    def updateInverseDepth(self, vel):
        B = self.lanes
        v = _f32(np.broadcast_to(np.asarray(vel, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        check(self.L.rbv_update_inverse_depth(self.h, ptr(v)))

    # fused per-frame sequences
    def track_g1(self, old, new, R):
        B = self.lanes
        R = _f32(np.broadcast_to(np.asarray(R, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        vel = np.zeros((B, 3), np.float32); Rv = np.zeros((B, 3, 3), np.float32)
        Wx = np.zeros((B, 6, 6), np.float32); X = np.zeros((B, 6), np.float32); ok = np.zeros(B, np.int32)
        check(self.L.rbv_track_g1(self.h, old.h, new.h, ptr(R), ptr(vel), ptr(Rv), ptr(Wx), ptr(X), ptr(ok)))
        return vel, Rv, Wx, X, ok

    def track_g2(self, old, new, R0, V, P_V, Rgva, max_radius=40.0, min_matches=0):
        B = self.lanes
        R0 = _f32(np.broadcast_to(np.asarray(R0, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        V = _f32(np.broadcast_to(np.asarray(V, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        PV = _f32(np.broadcast_to(np.asarray(P_V, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        Rg = _f32(np.broadcast_to(np.asarray(Rgva, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        n = np.zeros(B, np.int32); kf = np.zeros(B, np.int32)
        check(self.L.rbv_track_g2(self.h, old.h, new.h, ptr(R0), ptr(V), ptr(PV), ptr(Rg), max_radius, min_matches, ptr(n), ptr(kf)))
        return n, kf

    def track(self, old, new, R, max_radius=40.0, min_matches=0):
        """one tracker frame in one call (rbv_track: g1, host fusion, g2); R [B,3,3] = rotation since the previous frame"""
        B = self.lanes
        R = _f32(np.broadcast_to(np.asarray(R, np.float32).reshape(-1, 9), (B, 9)), (B, 9))
        Vg = np.zeros((B, 3), np.float32); PVg = np.zeros((B, 3, 3), np.float32)
        Wx = np.zeros((B, 6, 6), np.float32); X = np.zeros((B, 6), np.float32); ok = np.zeros(B, np.int32)
        V = np.zeros((B, 3), np.float32); n = np.zeros(B, np.int32); kf = np.zeros(B, np.int32)
        nk = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
        check(self.L.rbv_track(self.h, old.h, new.h, ptr(R), max_radius, min_matches, ptr(Vg), ptr(PVg), ptr(Wx), ptr(X), ptr(ok),
                               ptr(V), ptr(n), ptr(kf), ptr(nk), ptr(at)))
        return dict(Vg=Vg, P_Vg=PVg, Wx=Wx, X=X, ok=ok, V=V, matches=n, kf_matches=kf, keylines=nk, auto_threshold=at)

    # ---- IMU fusion block (rbv_fusion.cu; reference rebvio.cpp:146-165, 183-233, 255-273; core.cpp:264-414) ----
    def imu_reset(self, cfg=None):
        """state of a freshly constructed Rebvio for every lane; cfg = capi.ImuCfg (default: imu.hpp:153-168)"""
        if cfg is None:
            cfg = capi.ImuCfg(); self.L.rbv_default_imu_cfg(C.byref(cfg))
        self.imu_cfg = cfg
        check(self.L.rbv_imu_reset(self.h, C.byref(cfg)))

    def _imu_in(self, imu_in):
        a = np.ascontiguousarray(np.broadcast_to(np.asarray(imu_in, capi.IMU_IN).reshape(-1), (self.lanes,)))
        return a

    def imu_prerotate(self, imu_in):
        """rebvio.cpp:146-165 -> R.T() [B,3,3], the argument of the first rotateKeylines"""
        a = self._imu_in(imu_in)
        Rt = np.zeros((self.lanes, 3, 3), np.float32)
        check(self.L.rbv_imu_prerotate(self.h, a.ctypes.data, ptr(Rt)))
        return Rt

    def imu_fuse(self, Vg, Wx, X):
        """rebvio.cpp:183-233 + pose -> record array (capi.IMU_OUT) per lane"""
        B = self.lanes
        Vg = _f32(np.broadcast_to(np.asarray(Vg, np.float32).reshape(-1, 3), (B, 3)), (B, 3))
        Wx = _f32(np.broadcast_to(np.asarray(Wx, np.float32).reshape(-1, 36), (B, 36)), (B, 36))
        X = _f32(np.broadcast_to(np.asarray(X, np.float32).reshape(-1, 6), (B, 6)), (B, 6))
        out = np.zeros(B, capi.IMU_OUT)
        check(self.L.rbv_imu_fuse(self.h, ptr(Vg), ptr(Wx), ptr(X), out.ctypes.data))
        return out

    def track_imu(self, old, new, imu_in, max_radius=40.0, min_matches=0):
        """one tracker frame with the IMU filters, kernels only (rbv_track_imu)"""
        B = self.lanes
        a = self._imu_in(imu_in)
        Vg = np.zeros((B, 3), np.float32); PVg = np.zeros((B, 3, 3), np.float32)
        Wx = np.zeros((B, 6, 6), np.float32); X = np.zeros((B, 6), np.float32); ok = np.zeros(B, np.int32)
        n = np.zeros(B, np.int32); kf = np.zeros(B, np.int32); nk = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
        out = np.zeros(B, capi.IMU_OUT)
        check(self.L.rbv_track_imu(self.h, old.h, new.h, a.ctypes.data, max_radius, min_matches, ptr(Vg), ptr(PVg), ptr(Wx),
                                   ptr(X), ptr(ok), out.ctypes.data, ptr(n), ptr(kf), ptr(nk), ptr(at)))
        return dict(Vg=Vg, P_Vg=PVg, Wx=Wx, X=X, ok=ok, imu=out, V=out["V"].copy(), matches=n, kf_matches=kf, keylines=nk,
                    auto_threshold=at)

    # debug taps
    def read_image(self, which, lane=0):
        out = np.empty((self.cam.rows, self.cam.cols), np.float32)
        check(self.L.rbv_debug_read_image(self.h, lane, which, ptr(out)))
        return out

    def read_field(self, lane=0):
        i = np.empty((self.cam.rows, self.cam.cols), np.int32); d = np.empty_like(i)
        check(self.L.rbv_debug_read_field(self.h, lane, ptr(i), ptr(d)))
        return i, d

    def box_pass(self, scale, p, img):
        img = np.ascontiguousarray(img, np.float32)
        out = np.empty_like(img)
        check(self.L.rbv_debug_box_pass(self.h, scale, p, ptr(img), ptr(out)))
        return out

    def setup(self):
        pinv = np.zeros(75, np.float32); w = np.zeros(6, np.int32)
        check(self.L.rbv_debug_read_setup(self.h, ptr(pinv), ptr(w)))
        return pinv.reshape(3, 25), w

    def probe_count(self, lane=0):
        p = C.c_ulonglong(0)
        check(self.L.rbv_debug_read_probe_count(self.h, lane, C.byref(p)))
        return p.value
