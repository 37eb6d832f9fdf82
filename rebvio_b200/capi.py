"""ctypes binding of include/rebvio_b200.h. Loads rebvio_b200/librebvio_b200.so and fails loudly if it is missing
(build with `python -c "import __graft_entry__ as g; g.build()"` or `make -C rebvio_b200/csrc`)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RBV_LIB") or os.path.join(HERE, "librebvio_b200.so")   # RBV_LIB: another build of the same library (bisecting on the GPU box)

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")


class Camera(C.Structure):
    _fields_ = [("fm", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("rows", C.c_int), ("cols", C.c_int)]


class DetectorCfg(C.Structure):
    _fields_ = [("keylines_ref", C.c_int), ("keylines_max", C.c_int), ("pos_neg_threshold", C.c_float),
                ("dog_threshold", C.c_float), ("threshold", C.c_float), ("gain", C.c_float),
                ("max_threshold", C.c_float), ("min_threshold", C.c_float)]


class MapCfg(C.Structure):
    _fields_ = [("pixel_uncertainty_match", C.c_float), ("match_threshold_norm", C.c_float),
                ("match_threshold_angle", C.c_float), ("regularization_threshold", C.c_float)]


class ImuCfg(C.Structure):
    """types::ImuStateConfig (reference imu.hpp:153-168)"""
    _fields_ = [(n, C.c_float) for n in ("g_norm", "g_uncertainty", "g_norm_uncertainty", "acc_std_dev", "gyro_std_dev",
                                         "gyro_bias_std_dev", "vbias_std_dev", "scale_std_dev_mult", "scale_std_dev_max",
                                         "scale_stdd_dev_init")] + \
               [("init_bias", C.c_int), ("init_bias_frame_num", C.c_int), ("init_bias_guess", C.c_float * 3)]


# numpy mirrors of rbv_imu_in / rbv_imu_out (include/rebvio_b200.h): one record per lane
IMU_IN = np.dtype([("R", np.float32, 9), ("gyro", np.float32, 3), ("cacc", np.float32, 3), ("dt_s", np.float32),
                   ("frame_dt", np.float32)])
IMU_OUT = np.dtype([("Rt", np.float32, 9), ("R", np.float32, 9), ("R0", np.float32, 9), ("Rgva", np.float32, 9),
                    ("V", np.float32, 3), ("P_V", np.float32, 9), ("Xgv", np.float32, 6), ("Xgva", np.float32, 6),
                    ("Bg", np.float32, 3), ("Av", np.float32, 3), ("As", np.float32, 3), ("X", np.float32, 7),
                    ("orientation", np.float32, 3), ("position", np.float32, 3), ("K", np.float32),
                    ("initialized", np.int32), ("stopped", np.int32), ("num_frames", np.int32)])


class CoreCfg(C.Structure):
    _fields_ = [("search_range", C.c_float), ("reweight_distance", C.c_float), ("match_treshold", C.c_float),
                ("min_match_threshold", C.c_uint), ("iterations", C.c_uint),
                ("global_min_matches_threshold", C.c_uint), ("pixel_uncertainty", C.c_float),
                ("quantile_cutoff", C.c_float), ("quantile_num_bins", C.c_int), ("reshape_q_abs", C.c_float)]


KEYLINE_DTYPE = np.dtype([
    ("pos", np.float32, 2), ("pos_img", np.float32, 2), ("match_pos_img", np.float32, 2),
    ("gradient", np.float32, 2), ("match_gradient", np.float32, 2), ("gradient_norm", np.float32),
    ("match_gradient_norm", np.float32), ("rho", np.float32), ("sigma_rho", np.float32),
    ("id", np.int32), ("id_prev", np.int32), ("id_next", np.int32), ("match_id", np.int32),
    ("match_id_forward", np.int32), ("match_id_keyframe", np.int32), ("matches", np.uint32)])
assert KEYLINE_DTYPE.itemsize == 84

RBV_OK, RBV_ERR_ARG, RBV_ERR_CUDA, RBV_ERR_STATE = 0, 1, 2, 3
REDUCE_FAST, REDUCE_ORDERED = 0, 1
TUNE_FIELD, TUNE_BOX, TUNE_TRYVEL, TUNE_GRAPH, TUNE_FIELD_ROWS, TUNE_FE_STREAM, TUNE_FIELD_IN_FE, TUNE_ORDERED = 0, 1, 2, 3, 4, 5, 7, 8   # rbv_set_tuning keys (0 = automatic; see the header for the values)

vp = C.c_void_p
_SIG = {
    "rbv_default_camera": (None, [C.POINTER(Camera)]),
    "rbv_default_detector_cfg": (None, [C.POINTER(DetectorCfg)]),
    "rbv_default_map_cfg": (None, [C.POINTER(MapCfg)]),
    "rbv_default_core_cfg": (None, [C.POINTER(CoreCfg)]),
    "rbv_create": (C.c_int, [C.POINTER(Camera), C.POINTER(DetectorCfg), C.POINTER(MapCfg), C.POINTER(CoreCfg), C.c_int,
                            C.c_int, C.POINTER(vp)]),
    "rbv_destroy": (None, [vp]),
    "rbv_last_error": (C.c_char_p, []),
    "rbv_sync": (C.c_int, [vp]),
    "rbv_set_reduce_mode": (C.c_int, [vp, C.c_int]),
    "rbv_set_tuning": (C.c_int, [vp, C.c_int, C.c_int]),
    "rbv_lanes": (C.c_int, [vp]),
    "rbv_set_stream": (C.c_int, [vp, vp]),
    "rbv_launch_count": (C.c_ulonglong, [vp]),
    "rbv_kernel_count": (C.c_int, []),
    "rbv_kernel_name": (C.c_char_p, [C.c_int]),
    "rbv_box_pass_noalloc": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp]),
    "rbv_profile_set": (C.c_int, [vp, C.c_int]),
    "rbv_profile_reset": (C.c_int, [vp]),
    "rbv_profile_read": (C.c_int, [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
    "rbv_map_create": (C.c_int, [vp, C.POINTER(vp)]),
    "rbv_map_release": (None, [vp, vp]),
    "rbv_detect": (C.c_int, [vp, vp, C.POINTER(vp), C.c_size_t, vp, vp, vp, vp]),
    "rbv_detect_u8": (C.c_int, [vp, vp, C.POINTER(vp), C.c_size_t, vp, vp, vp, vp]),
    "rbv_prefetch_frames": (C.c_int, [vp, C.POINTER(vp), C.c_size_t, C.c_int]),
    "rbv_detect_wait": (C.c_int, [vp, vp, vp, vp, vp]),
    "rbv_detect_device": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "rbv_servo_threshold": (C.c_float, [C.POINTER(DetectorCfg), C.c_float, C.c_int]),
    "rbv_build_distance_field": (C.c_int, [vp, vp]),
    "rbv_rotate_keylines": (C.c_int, [vp, vp, vp]),
    "rbv_estimate_quantile": (C.c_int, [vp, vp, C.c_float, C.c_int, vp]),
    "rbv_reset_residuals": (C.c_int, [vp, vp]),
    "rbv_try_vel": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "rbv_minimize_vel": (C.c_int, [vp, vp, vp, vp, vp]),
    "rbv_forward_match": (C.c_int, [vp, vp, vp, vp]),
    "rbv_ext_rot_vel": (C.c_int, [vp, vp, vp, vp, vp, vp]),
    "rbv_directed_match": (C.c_int, [vp, vp, vp, vp, vp, vp, C.c_float, vp, vp]),
    "rbv_regularize": (C.c_int, [vp, vp, vp]),
    "rbv_update_inverse_depth": (C.c_int, [vp, vp]),
    "rbv_track_g1": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rbv_track_g2": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp]),
    "rbv_set_undistort": (C.c_int, [vp, vp, vp]),
    "rbv_visual_fusion": (C.c_int, [C.c_int, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rbv_track": (C.c_int, [vp, vp, vp, vp, C.c_float, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rbv_default_imu_cfg": (None, [vp]),
    "rbv_imu_reset": (C.c_int, [vp, vp]),
    "rbv_imu_prerotate": (C.c_int, [vp, vp, vp]),
    "rbv_imu_fuse": (C.c_int, [vp, vp, vp, vp, vp]),
    "rbv_track_imu": (C.c_int, [vp, vp, vp, vp, C.c_float, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rbv_map_size": (C.c_int, [vp, vp, vp]),
    "rbv_map_threshold": (C.c_int, [vp, vp, vp]),
    "rbv_map_set_threshold": (C.c_int, [vp, vp, vp]),
    "rbv_map_download": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_int, vp]),
    "rbv_map_upload": (C.c_int, [vp, vp, C.c_int, vp, vp, C.c_int]),
    "rbv_debug_read_image": (C.c_int, [vp, C.c_int, C.c_int, vp]),
    "rbv_debug_read_mask": (C.c_int, [vp, vp, C.c_int, vp]),
    "rbv_debug_read_field": (C.c_int, [vp, C.c_int, vp, vp]),
    "rbv_debug_read_residuals": (C.c_int, [vp, vp, C.c_int, vp, C.c_int]),
    "rbv_debug_box_pass": (C.c_int, [vp, C.c_int, C.c_int, vp, vp]),
    "rbv_debug_read_setup": (C.c_int, [vp, vp, vp]),
    "rbv_debug_read_lm_trace": (C.c_int, [vp, C.c_int, vp, C.c_int]),
    "rbv_debug_read_probe_count": (C.c_int, [vp, C.c_int, vp]),
    "rbv_debug_read_addends": (C.c_int, [vp, C.c_int, C.c_int, vp, C.c_int]),
}

EXPORTS = sorted(_SIG)
_lib = None


def load():
    """dlopen the CUDA library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build the CUDA extension first (__graft_entry__.build()); "
                               "rebvio_b200 has no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIG.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


class RbvError(RuntimeError):
    pass


def check(rc):
    if rc != RBV_OK:
        raise RbvError(f"rbv error {rc}: {load().rbv_last_error().decode(errors='replace')}")


def ptr(a):
    return None if a is None else a.ctypes.data_as(vp)


def defaults():
    L = load()
    cam, det, mp, core = Camera(), DetectorCfg(), MapCfg(), CoreCfg()
    L.rbv_default_camera(C.byref(cam)); L.rbv_default_detector_cfg(C.byref(det))
    L.rbv_default_map_cfg(C.byref(mp)); L.rbv_default_core_cfg(C.byref(core))
    return cam, det, mp, core
