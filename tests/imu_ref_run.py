"""Runs the UNMODIFIED reference (oracle/_ref) through its deterministic single-threaded pipeline with IMU samples and
returns the per-frame records. Executed in a fresh process: Core::estimateLs4Acceleration / estimateMeanAcceleration
keep their filter state in function-local statics (core.cpp:292-298, 340-343), so one process = one sequence."""
from __future__ import annotations

import numpy as np

FIELDS = ("orientation", "position", "V", "Vg", "Xv", "Xgv", "Xgva", "P_Vg", "P_V", "R", "Rgva", "R0", "W_Xv", "imu_R",
          "imu_gyro", "imu_cacc", "Bg", "Av", "As", "sabX", "Rt1")
SCALARS = ("score", "K", "frame_dt", "imu_dt_s", "n_new", "n_old", "fwd_matches", "klm_num", "kf_matches", "reg_num", "ok",
           "estimated", "stopped", "initialized", "ts_us")


def run(name, n_frames, q=None):
    from oracle import pyref
    from rebvio_b200 import synth
    import common
    spec = synth.config(name, n_frames=n_frames)
    seq = synth.generate(spec)
    cfg = common.cfgs_from_spec(spec)
    ref = pyref.RefLib(cfg[0], cfg[1], cfg[3])
    recs, j = [], 0
    for k in range(n_frames):
        while j < len(seq.imu_ts_us) and seq.imu_ts_us[j] <= seq.ts_us[k]:
            ref.imu(seq.imu_ts_us[j], seq.gyro[j], seq.acc[j]); j += 1
        o = ref.frame(synth.to_float3(seq.frames[k]), seq.ts_us[k])
        r = {f: np.array(getattr(o, f), np.float32) for f in FIELDS}
        r.update({f: getattr(o, f) for f in SCALARS})
        recs.append(r)
    Rc2i, tc2i = ref.camera_extrinsics()
    recs[0]["Rc2i"], recs[0]["tc2i"] = Rc2i, tc2i
    ref.close()
    if q is not None:
        q.put(recs)
    return recs


def run_in_subprocess(name, n_frames):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    p = ctx.Process(target=run, args=(name, n_frames, q))
    p.start()
    recs = q.get(timeout=600)
    p.join(60)
    return recs
