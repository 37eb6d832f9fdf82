"""ctypes bindings of the parity checkers. TEST INFRASTRUCTURE ONLY.

  RefLib   -> oracle/_ref/librebvio_ref.so      : the UNMODIFIED reference compiled from /root/reference
                                                  (oracle/ref_harness.cpp, oracle/Makefile target `ref`)
  RefLib(fast=True) -> librebvio_ref_fast.so    : same sources, -O3 + FMA; timing only
  OracleLib -> oracle/librebvio_oracle.so       : plain-C restatement (oracle/rebvio_oracle.c)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
Nothing under rebvio_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


class Camera(C.Structure):
    _fields_ = [("fm", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("rows", C.c_int), ("cols", C.c_int)]


class DetectorCfg(C.Structure):
    _fields_ = [("keylines_ref", C.c_int), ("keylines_max", C.c_int), ("pos_neg_threshold", C.c_float),
                ("dog_threshold", C.c_float), ("threshold", C.c_float), ("gain", C.c_float),
                ("max_threshold", C.c_float), ("min_threshold", C.c_float)]


class MapCfg(C.Structure):
    _fields_ = [("pixel_uncertainty_match", C.c_float), ("match_threshold_norm", C.c_float),
                ("match_threshold_angle", C.c_float), ("regularization_threshold", C.c_float)]


class CoreCfg(C.Structure):
    _fields_ = [("search_range", C.c_float), ("reweight_distance", C.c_float), ("match_treshold", C.c_float),
                ("min_match_threshold", C.c_uint), ("iterations", C.c_uint),
                ("global_min_matches_threshold", C.c_uint), ("pixel_uncertainty", C.c_float),
                ("quantile_cutoff", C.c_float), ("quantile_num_bins", C.c_int), ("reshape_q_abs", C.c_float)]


# numpy mirror of rbv_keyline / rebvio::types::KeyLine (84 bytes, keyline.hpp:24-60)
KEYLINE_DTYPE = np.dtype([
    ("pos", np.float32, 2), ("pos_img", np.float32, 2), ("match_pos_img", np.float32, 2),
    ("gradient", np.float32, 2), ("match_gradient", np.float32, 2), ("gradient_norm", np.float32),
    ("match_gradient_norm", np.float32), ("rho", np.float32), ("sigma_rho", np.float32),
    ("id", np.int32), ("id_prev", np.int32), ("id_next", np.int32), ("match_id", np.int32),
    ("match_id_forward", np.int32), ("match_id_keyframe", np.int32), ("matches", np.uint32)])
assert KEYLINE_DTYPE.itemsize == 84


class FrameOut(C.Structure):
    """ref_frame_out of oracle/ref_harness.cpp"""
    _fields_ = [("ts_us", C.c_uint64), ("orientation", C.c_float * 3), ("position", C.c_float * 3),
                ("V", C.c_float * 3), ("Vg", C.c_float * 3), ("Xv", C.c_float * 6), ("Xgv", C.c_float * 6),
                ("Xgva", C.c_float * 6), ("P_Vg", C.c_float * 9), ("P_V", C.c_float * 9), ("R", C.c_float * 9),
                ("Rgva", C.c_float * 9), ("R0", C.c_float * 9), ("W_Xv", C.c_float * 36), ("score", C.c_float),
                ("K", C.c_float), ("frame_dt", C.c_float), ("n_new", C.c_int), ("n_old", C.c_int),
                ("fwd_matches", C.c_int), ("klm_num", C.c_int), ("kf_matches", C.c_int), ("reg_num", C.c_int),
                ("ok", C.c_int), ("estimated", C.c_int), ("stopped", C.c_int), ("old_map", C.c_int),
                ("new_map", C.c_int), ("imu_R", C.c_float * 9), ("imu_gyro", C.c_float * 3), ("imu_cacc", C.c_float * 3),
                ("imu_dt_s", C.c_float), ("Bg", C.c_float * 3), ("Av", C.c_float * 3), ("As", C.c_float * 3),
                ("sabX", C.c_float * 7), ("Rt1", C.c_float * 9), ("initialized", C.c_int)]


def default_cfgs(cam=None, det=None, core=None, mapc=None):
    """Defaults of the reference structs (camera.hpp:26-37, edge_detector.hpp:19-32, edge_map.hpp:19-26,
    core.hpp:82-95), stated independently of the product's rbv_default_* (tests compare the two)."""
    c = Camera(np.float32(0.5) * (np.float32(458.654) + np.float32(457.296)), 367.215, 248.375, 480, 752)
    d = DetectorCfg(12000, 16000, 0.4, 0.095259868922420, 0.01, 5e-7, 0.5, 0.005)
    m = MapCfg(2.0, 1.0, 45.0, 0.5)
    k = CoreCfg(40.0, 2.0, 0.5, 0, 5, 500, 1.0, 0.9, 100, 1e-4)
    for obj, over in ((c, cam), (d, det), (k, core), (m, mapc)):
        for key, val in (over or {}).items():
            setattr(obj, key, val)
    return c, d, m, k


def available(fast: bool = False) -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "librebvio_ref_fast.so" if fast else "librebvio_ref.so"))


class RefLib:
    """One reference pipeline instance (EdgeDetector + Core + deterministic driver)."""

    def __init__(self, cam: Camera, det: DetectorCfg, core: CoreCfg, fast: bool = False):
        path = os.path.join(HERE, "_ref", "librebvio_ref_fast.so" if fast else "librebvio_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
        L = self.L = C.CDLL(path)
        self.rows, self.cols, self.kmax = cam.rows, cam.cols, det.keylines_max
        self.n = cam.rows * cam.cols
        vp = C.c_void_p
        L.ref_create.restype = vp
        L.ref_create.argtypes = [C.POINTER(Camera), C.POINTER(DetectorCfg), C.POINTER(CoreCfg)]
        sig = {
            "ref_destroy": (None, [vp]),
            "ref_scale_setup": (None, [vp, i32p, f32p]),
            "ref_divisors": (None, [vp, C.c_int, C.c_int, f32p]),
            "ref_integral": (None, [vp, f32p, f32p]),
            "ref_box_pass": (None, [vp, C.c_int, C.c_int, f32p, f32p]),
            "ref_scale_space": (None, [vp, f32p, f32p, f32p, f32p, f32p]),
            "ref_get_threshold": (C.c_float, [vp]),
            "ref_set_threshold": (None, [vp, C.c_float]),
            "ref_set_last_count": (None, [vp, C.c_int]),
            "ref_detect": (C.c_int, [vp, f32p, C.c_uint64]),
            "ref_detector_mask": (None, [vp, i32p]),
            "ref_map_free": (None, [vp, C.c_int]),
            "ref_map_new": (C.c_int, [vp, C.c_uint64]),
            "ref_map_size": (C.c_int, [vp, C.c_int]),
            "ref_map_threshold": (C.c_float, [vp, C.c_int]),
            "ref_map_set_threshold": (None, [vp, C.c_int, C.c_float]),
            "ref_map_get": (None, [vp, C.c_int, C.c_void_p]),
            "ref_map_set": (None, [vp, C.c_int, C.c_void_p, C.c_void_p, C.c_int]),
            "ref_map_pixels": (None, [vp, C.c_int, u32p]),
            "ref_map_mask_size": (C.c_int, [vp, C.c_int]),
            "ref_build_distance_field": (None, [vp, C.c_int]),
            "ref_field": (None, [vp, i32p, i32p]),
            "ref_rotate": (None, [vp, C.c_int, f32p]),
            "ref_quantile": (C.c_float, [vp, C.c_int, C.c_float, C.c_int]),
            "ref_try_vel": (C.c_float, [vp, C.c_int, f32p, C.c_float, f32p, f32p, f32p]),
            "ref_minimize_vel": (C.c_float, [vp, C.c_int, f32p, f32p]),
            "ref_minimize_vel_trace": (C.c_float, [vp, C.c_int, f32p, f32p, f32p, C.c_void_p]),
            "ref_forward_match": (C.c_int, [vp, C.c_int, C.c_int]),
            "ref_ext_rot_vel": (C.c_int, [vp, C.c_int, f32p, f32p, f32p]),
            "ref_directed_match": (C.c_int, [vp, C.c_int, C.c_int, f32p, f32p, f32p, C.c_float, C.POINTER(C.c_int)]),
            "ref_regularize": (C.c_int, [vp, C.c_int]),
            "ref_update_inverse_depth": (None, [vp, f32p]),
            "ref_pipeline_imu": (None, [vp, C.c_uint64, f32p, f32p]),
            "ref_pipeline_frame": (C.c_int, [vp, f32p, C.c_uint64, C.POINTER(FrameOut)]),
            "ref_timers": (None, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double * 9)]),
            "ref_timers_reset": (None, [vp]),
            "ref_camera_extrinsics": (None, [f32p, f32p]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        self.h = L.ref_create(C.byref(cam), C.byref(det), C.byref(core))

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    # ---- scale space ----
    def scale_setup(self):
        w = np.zeros(6, np.int32); s = np.zeros(2, np.float32)
        self.L.ref_scale_setup(self.h, w, s)
        return w, s

    def divisors(self, scale, p):
        out = np.empty((self.rows, self.cols), np.float32)
        self.L.ref_divisors(self.h, scale, p, out)
        return out

    def integral(self, img):
        out = np.empty((self.rows, self.cols), np.float32)
        self.L.ref_integral(self.h, np.ascontiguousarray(img, np.float32), out)
        return out

    def box_pass(self, scale, p, img):
        out = np.empty((self.rows, self.cols), np.float32)
        self.L.ref_box_pass(self.h, scale, p, np.ascontiguousarray(img, np.float32), out)
        return out

    def scale_space(self, img):
        o = [np.empty((self.rows, self.cols), np.float32) for _ in range(4)]
        self.L.ref_scale_space(self.h, np.ascontiguousarray(img, np.float32), *o)
        return o  # s0, s1, dog, mag

    # ---- front end ----
    def get_threshold(self): return self.L.ref_get_threshold(self.h)
    def set_threshold(self, t): self.L.ref_set_threshold(self.h, t)
    def set_last_count(self, n): self.L.ref_set_last_count(self.h, n)

    def detect(self, img, ts_us=0):
        return self.L.ref_detect(self.h, np.ascontiguousarray(img, np.float32), ts_us)

    def detector_mask(self):
        out = np.empty((self.rows, self.cols), np.int32)
        self.L.ref_detector_mask(self.h, out)
        return out

    def map_free(self, m): self.L.ref_map_free(self.h, m)
    def map_new(self, ts_us=0): return self.L.ref_map_new(self.h, ts_us)
    def map_size(self, m): return self.L.ref_map_size(self.h, m)
    def map_threshold(self, m): return self.L.ref_map_threshold(self.h, m)
    def map_set_threshold(self, m, t): self.L.ref_map_set_threshold(self.h, m, t)

    def map_get(self, m):
        kl = np.zeros(self.map_size(m), KEYLINE_DTYPE)
        if len(kl):
            self.L.ref_map_get(self.h, m, kl.ctypes.data)
        return kl

    def map_set(self, m, kl, pixel_idx=None):
        kl = np.ascontiguousarray(kl, KEYLINE_DTYPE)
        px = None if pixel_idx is None else np.ascontiguousarray(pixel_idx, np.uint32)
        self.L.ref_map_set(self.h, m, kl.ctypes.data, None if px is None else px.ctypes.data, len(kl))

    def map_pixels(self, m):
        out = np.zeros(max(1, self.map_size(m)), np.uint32)
        self.L.ref_map_pixels(self.h, m, out)
        return out[:self.map_size(m)]

    # ---- tracker ----
    def build_distance_field(self, m): self.L.ref_build_distance_field(self.h, m)

    def field(self):
        i = np.empty((self.rows, self.cols), np.int32); d = np.empty((self.rows, self.cols), np.int32)
        self.L.ref_field(self.h, i, d)
        return i, d

    def rotate(self, m, R): self.L.ref_rotate(self.h, m, np.ascontiguousarray(R, np.float32).reshape(9))
    def quantile(self, m, pct=0.9, bins=100): return self.L.ref_quantile(self.h, m, pct, bins)

    def try_vel(self, m, vel, sigma_rho_min, residuals):
        JtJ = np.zeros(9, np.float32); JtF = np.zeros(3, np.float32)
        s = self.L.ref_try_vel(self.h, m, np.ascontiguousarray(vel, np.float32), sigma_rho_min, residuals, JtJ, JtF)
        return s, JtJ.reshape(3, 3), JtF

    def minimize_vel(self, m, vel):
        v = np.array(vel, np.float32); Rv = np.zeros(9, np.float32)
        s = self.L.ref_minimize_vel(self.h, m, v, Rv)
        return s, v, Rv.reshape(3, 3)

    def minimize_vel_trace(self, m, vel, iterations=5, want_residuals=True):
        v = np.array(vel, np.float32); Rv = np.zeros(9, np.float32)
        tr = np.zeros((iterations + 1, 16), np.float32)
        res = np.zeros(max(1, self.map_size(m)), np.float32)
        s = self.L.ref_minimize_vel_trace(self.h, m, v, Rv, tr.reshape(-1), res.ctypes.data if want_residuals else None)
        return s, v, Rv.reshape(3, 3), tr, res[:self.map_size(m)]

    def forward_match(self, old_m, new_m): return self.L.ref_forward_match(self.h, old_m, new_m)

    def ext_rot_vel(self, new_m, vel):
        Wx = np.zeros(36, np.float32); X = np.zeros(6, np.float32)
        ok = self.L.ref_ext_rot_vel(self.h, new_m, np.ascontiguousarray(vel, np.float32), Wx, X)
        return ok, Wx.reshape(6, 6), X

    def directed_match(self, new_m, old_m, vel, Rvel, Rback, max_radius=40.0):
        kf = C.c_int(0)
        n = self.L.ref_directed_match(self.h, new_m, old_m, np.ascontiguousarray(vel, np.float32),
                                      np.ascontiguousarray(Rvel, np.float32).reshape(9),
                                      np.ascontiguousarray(Rback, np.float32).reshape(9), max_radius, C.byref(kf))
        return n, kf.value

    def regularize(self, m): return self.L.ref_regularize(self.h, m)
    def update_inverse_depth(self, vel): self.L.ref_update_inverse_depth(self.h, np.ascontiguousarray(vel, np.float32))

    # ---- pipeline ----
    def imu(self, ts_us, gyro, acc):
        self.L.ref_pipeline_imu(self.h, int(ts_us), np.ascontiguousarray(gyro, np.float32),
                                np.ascontiguousarray(acc, np.float32))

    def frame(self, img, ts_us) -> FrameOut:
        out = FrameOut()
        self.L.ref_pipeline_frame(self.h, np.ascontiguousarray(img, np.float32), int(ts_us), C.byref(out))
        return out

    def timers(self):
        d, t, st = C.c_double(), C.c_double(), (C.c_double * 9)()
        self.L.ref_timers(self.h, C.byref(d), C.byref(t), C.byref(st))
        return d.value, t.value, list(st)

    def timers_reset(self): self.L.ref_timers_reset(self.h)

    def camera_extrinsics(self):
        """Camera::getRc2i / getTc2i of a default-constructed reference camera"""
        R, t = np.zeros(9, np.float32), np.zeros(3, np.float32)
        self.L.ref_camera_extrinsics(R, t)
        return R.reshape(3, 3), t


class OracleLib:
    """Plain-C restatement (oracle/rebvio_oracle.c). Same method names as RefLib; maps are opaque handles."""

    def __init__(self, cam: Camera, det: DetectorCfg, core: CoreCfg, mapc: MapCfg | None = None):
        path = os.path.join(HERE, "librebvio_oracle.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle oracle`")
        L = self.L = C.CDLL(path)
        self.cam, self.det, self.core = cam, det, core
        self.mapc = mapc or default_cfgs()[2]
        self.rows, self.cols, self.kmax = cam.rows, cam.cols, det.keylines_max
        vp = C.c_void_p
        L.orc_create.restype = vp
        L.orc_create.argtypes = [C.POINTER(Camera), C.POINTER(DetectorCfg), C.POINTER(MapCfg), C.POINTER(CoreCfg)]
        sig = {
            "orc_destroy": (None, [vp]),
            "orc_scale_setup": (None, [vp, i32p, f32p]),
            "orc_divisors": (None, [vp, C.c_int, C.c_int, f32p]),
            "orc_integral": (None, [vp, f32p, f32p]),
            "orc_box_pass": (None, [vp, C.c_int, C.c_int, f32p, f32p]),
            "orc_scale_space": (None, [vp, f32p, f32p, f32p, f32p, f32p]),
            "orc_pinv": (None, [vp, f32p]),
            "orc_map_new": (vp, [vp]),
            "orc_map_free": (None, [vp]),
            "orc_map_size": (C.c_int, [vp]),
            "orc_map_threshold": (C.c_float, [vp]),
            "orc_map_set_threshold": (None, [vp, C.c_float]),
            "orc_map_get": (None, [vp, C.c_void_p]),
            "orc_map_set": (None, [vp, C.c_void_p, C.c_void_p, C.c_int]),
            "orc_map_pixels": (None, [vp, u32p]),
            "orc_servo_threshold": (C.c_float, [C.POINTER(DetectorCfg), C.c_float, C.c_int]),
            "orc_detect": (C.c_int, [vp, f32p, C.c_float, vp]),
            "orc_detector_mask": (None, [vp, i32p]),
            "orc_build_distance_field": (None, [vp, vp]),
            "orc_field": (None, [vp, i32p, i32p]),
            "orc_rotate": (None, [vp, vp, f32p]),
            "orc_quantile": (C.c_float, [vp, C.c_float, C.c_int]),
            "orc_try_vel": (C.c_float, [vp, vp, f32p, C.c_float, f32p, f32p, f32p]),
            "orc_minimize_vel": (C.c_float, [vp, vp, f32p, f32p, C.c_void_p, C.c_void_p]),
            "orc_forward_match": (C.c_int, [vp, vp]),
            "orc_ext_rot_vel": (C.c_int, [vp, vp, f32p, f32p, f32p, C.c_void_p]),
            "orc_directed_match": (C.c_int, [vp, vp, vp, f32p, f32p, f32p, C.c_float, C.POINTER(C.c_int),
                                            C.POINTER(C.c_ulonglong)]),
            "orc_regularize": (C.c_int, [vp, vp]),
            "orc_update_inverse_depth": (None, [vp, f32p]),
            "orc_solve6": (None, [f32p, f32p, f32p]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        self.h = L.orc_create(C.byref(cam), C.byref(det), C.byref(self.mapc), C.byref(core))
        self.threshold = det.threshold   # config_->threshold (mutated by the servo)
        self.last_count = 0              # EdgeDetector::keylines_count_
        self.last_probes = 0

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h)
            self.h = None

    def scale_setup(self):
        w = np.zeros(6, np.int32); s = np.zeros(2, np.float32)
        self.L.orc_scale_setup(self.h, w, s)
        return w, s

    def _img(self): return np.empty((self.rows, self.cols), np.float32)

    def divisors(self, scale, p):
        out = self._img(); self.L.orc_divisors(self.h, scale, p, out); return out

    def integral(self, img):
        out = self._img(); self.L.orc_integral(self.h, np.ascontiguousarray(img, np.float32), out); return out

    def box_pass(self, scale, p, img):
        out = self._img(); self.L.orc_box_pass(self.h, scale, p, np.ascontiguousarray(img, np.float32), out); return out

    def scale_space(self, img):
        o = [self._img() for _ in range(4)]
        self.L.orc_scale_space(self.h, np.ascontiguousarray(img, np.float32), *o)
        return o

    def pinv(self):
        out = np.zeros(75, np.float32); self.L.orc_pinv(self.h, out); return out.reshape(3, 25)

    # maps
    def map_new(self, ts_us=0): return self.L.orc_map_new(self.h)
    def map_free(self, m): self.L.orc_map_free(m)
    def map_size(self, m): return self.L.orc_map_size(m)
    def map_threshold(self, m): return self.L.orc_map_threshold(m)
    def map_set_threshold(self, m, t): self.L.orc_map_set_threshold(m, t)

    def map_get(self, m):
        kl = np.zeros(self.map_size(m), KEYLINE_DTYPE)
        if len(kl):
            self.L.orc_map_get(m, kl.ctypes.data)
        return kl

    def map_set(self, m, kl, pixel_idx=None):
        kl = np.ascontiguousarray(kl, KEYLINE_DTYPE)
        px = None if pixel_idx is None else np.ascontiguousarray(pixel_idx, np.uint32)
        self.L.orc_map_set(m, kl.ctypes.data, None if px is None else px.ctypes.data, len(kl))

    def map_pixels(self, m):
        out = np.zeros(max(1, self.map_size(m)), np.uint32)
        self.L.orc_map_pixels(m, out)
        return out[:self.map_size(m)]

    # front end
    def get_threshold(self): return self.threshold
    def set_threshold(self, t): self.threshold = float(np.float32(t))
    def set_last_count(self, n): self.last_count = n

    def detect(self, img, ts_us=0):
        """EdgeDetector::detect incl. the servo (edge_detector.cpp:30-43)."""
        self.threshold = self.L.orc_servo_threshold(C.byref(self.det), self.threshold, self.last_count)
        m = self.map_new()
        self.last_count = self.L.orc_detect(self.h, np.ascontiguousarray(img, np.float32), self.threshold, m)
        return m

    def detector_mask(self):
        out = np.empty((self.rows, self.cols), np.int32); self.L.orc_detector_mask(self.h, out); return out

    # tracker
    def build_distance_field(self, m): self.L.orc_build_distance_field(self.h, m)

    def field(self):
        i = np.empty((self.rows, self.cols), np.int32); d = np.empty((self.rows, self.cols), np.int32)
        self.L.orc_field(self.h, i, d)
        return i, d

    def rotate(self, m, R): self.L.orc_rotate(self.h, m, np.ascontiguousarray(R, np.float32).reshape(9))
    def quantile(self, m, pct=0.9, bins=100): return self.L.orc_quantile(m, pct, bins)

    def try_vel(self, m, vel, sigma_rho_min, residuals):
        JtJ = np.zeros(9, np.float32); JtF = np.zeros(3, np.float32)
        s = self.L.orc_try_vel(self.h, m, np.ascontiguousarray(vel, np.float32), sigma_rho_min, residuals, JtJ, JtF)
        return s, JtJ.reshape(3, 3), JtF

    def minimize_vel(self, m, vel):
        s, v, Rv, _, _ = self.minimize_vel_trace(m, vel)
        return s, v, Rv

    def minimize_vel_trace(self, m, vel, iterations=None, want_residuals=True):
        it = self.core.iterations if iterations is None else iterations
        v = np.array(vel, np.float32); Rv = np.zeros(9, np.float32)
        tr = np.zeros((it + 1, 16), np.float32)
        res = np.zeros(max(1, self.map_size(m)), np.float32)
        s = self.L.orc_minimize_vel(self.h, m, v, Rv, tr.ctypes.data, res.ctypes.data)
        return s, v, Rv.reshape(3, 3), tr, res[:self.map_size(m)]

    def forward_match(self, old_m, new_m): return self.L.orc_forward_match(old_m, new_m)

    def ext_rot_vel(self, new_m, vel, want_jtf=False):
        Wx = np.zeros(36, np.float32); X = np.zeros(6, np.float32); JtF = np.zeros(6, np.float32)
        ok = self.L.orc_ext_rot_vel(self.h, new_m, np.ascontiguousarray(vel, np.float32), Wx, X, JtF.ctypes.data)
        if want_jtf:
            return ok, Wx.reshape(6, 6), X, JtF
        return ok, Wx.reshape(6, 6), X

    def directed_match(self, new_m, old_m, vel, Rvel, Rback, max_radius=40.0):
        kf = C.c_int(0); pr = C.c_ulonglong(0)
        n = self.L.orc_directed_match(self.h, new_m, old_m, np.ascontiguousarray(vel, np.float32),
                                      np.ascontiguousarray(Rvel, np.float32).reshape(9),
                                      np.ascontiguousarray(Rback, np.float32).reshape(9), max_radius, C.byref(kf),
                                      C.byref(pr))
        self.last_probes = pr.value
        return n, kf.value

    def regularize(self, m): return self.L.orc_regularize(self.h, m)
    def update_inverse_depth(self, vel): self.L.orc_update_inverse_depth(self.h, np.ascontiguousarray(vel, np.float32))

    def solve6(self, A, b):
        x = np.zeros(6, np.float32)
        self.L.orc_solve6(np.ascontiguousarray(A, np.float32).reshape(36), np.ascontiguousarray(b, np.float32), x)
        return x
