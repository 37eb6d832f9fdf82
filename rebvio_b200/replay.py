"""Headless replay: a recorded (or synthetic) camera + IMU sequence in, the reference's odometry log out.

Stands in for the ROS node of the reference (ros_rebvio/src/ros_rebvio.cpp:21-50 subscribes image + IMU topics and
forwards them to Rebvio::imageCallback / imuCallback; :89-126 publishes the odometry) with files instead of topics, in
the deterministic order of the reference's own replay test (ros_rebvio/test/test_ros_rebvio.cpp:29-45: every IMU sample
up to an image's time stamp is delivered before that image). Per frame: front end (rbv_detect_u8), then one
rbv_track_imu call - prerotate, tracker first half, IMU fusion kernel, second half - and one odometry line
(`ts rx ry rz px py pz`, 6 decimals; rebvio.cpp:270-286).

    python -m rebvio_b200.replay --synthetic C5:320x240 --frames 60 --out odometry.txt
    python -m rebvio_b200.replay --npz sequence.npz --out odometry.txt
        (npz: frames uint8 [n,rows,cols] already undistorted, ts_us uint64 [n], imu_ts_us uint64 [m], gyro / acc float32 [m,3],
         optional cam = (fm, cx, cy), keylines = (ref, max))

There is no CPU fallback: the library must load and a B200 must be present.
"""
from __future__ import annotations

import argparse
import sys

import numpy as np

from . import capi, imu
from .pipeline import BatchedPipeline


def replay(frames_u8, ts_us, imu_ts_us, gyro, acc, cam, det, mapc, core, imu_cfg=True, device=0, out=None):
    """-> list of (ts_us, orientation[3], position[3], record) for every estimated frame; stops like the reference when a
    lane reports NaN or too few matches (rebvio.cpp:236-252)"""
    pipe = BatchedPipeline(cam, det, mapc, core, lanes=1, device=device, imu_cfg=imu_cfg)
    res, j, prev_ts = [], 0, None
    try:
        for k in range(len(frames_u8)):
            ts = int(ts_us[k])
            integ = imu.IntegratedImu()                                # one integrator per edge map (edge_map.hpp: imu_ member)
            while j < len(imu_ts_us) and int(imu_ts_us[j]) <= ts:      # rebvio.cpp:70-80
                integ.add(imu_ts_us[j], gyro[j], acc[j]); j += 1
            rec = None                                                 # the first map's integral is never read (it is only ever the old map)
            if prev_ts is not None:
                rec = integ.get(frame_dt=np.float32(np.float32(ts - prev_ts) / 1000000.0))
            o = pipe.push(np.ascontiguousarray(frames_u8[k][None]), ts_us=ts, imu_in=rec)
            prev_ts = ts
            if rec is None:
                continue
            r = o["imu"][0]
            line = imu.odometry_line(ts, r["orientation"], r["position"])
            if out is not None:
                out.write(line + "\n")
            res.append((ts, r["orientation"].copy(), r["position"].copy(), r.copy(), int(o["matches"][0])))
            if int(r["stopped"]) or int(o["matches"][0]) < int(core.global_min_matches_threshold):
                break
    finally:
        pipe.close()
    return res


def write_sequence(path, frames_u8, ts_us, imu_ts_us, gyro, acc, cam, keylines_ref=0, keylines_max=0):
    """the sequence file of the C++ replay (rebvio_b200/cpp/rebvio_replay.cpp, which drives the rebvio_b200::Rebvio façade)"""
    import struct
    frames_u8 = np.ascontiguousarray(frames_u8, np.uint8)
    n, rows, cols = frames_u8.shape
    with open(path, "wb") as f:
        f.write(b"RBVSEQ1\0")
        f.write(struct.pack("<6i3f", rows, cols, n, len(imu_ts_us), int(keylines_ref), int(keylines_max), cam.fm, cam.cx, cam.cy))
        f.write(np.ascontiguousarray(ts_us, "<u8").tobytes())
        f.write(frames_u8.tobytes())
        f.write(np.ascontiguousarray(imu_ts_us, "<u8").tobytes())
        f.write(np.ascontiguousarray(gyro, "<f4").tobytes())
        f.write(np.ascontiguousarray(acc, "<f4").tobytes())


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--synthetic", default="", help="sequence name of rebvio_b200.synth (e.g. C2, C5:320x240)")
    ap.add_argument("--npz", default="")
    ap.add_argument("--frames", type=int, default=60)
    ap.add_argument("--out", default="-")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    cam, det, mapc, core = capi.defaults()
    if a.synthetic:
        from . import synth
        spec = synth.config(a.synthetic, n_frames=a.frames)
        seq = synth.generate(spec)
        frames, ts, its, gyro, acc = seq.frames, seq.ts_us, seq.imu_ts_us, seq.gyro, seq.acc
        cam.fm, cam.cx, cam.cy = spec.cam.fm, spec.cam.cx, spec.cam.cy
        det.keylines_ref, det.keylines_max = spec.keylines_ref, spec.keylines_max
    elif a.npz:
        z = np.load(a.npz)
        frames, ts, its, gyro, acc = z["frames"], z["ts_us"], z["imu_ts_us"], z["gyro"], z["acc"]
        if "cam" in z: cam.fm, cam.cx, cam.cy = [float(x) for x in z["cam"]]
        if "keylines" in z: det.keylines_ref, det.keylines_max = [int(x) for x in z["keylines"]]
    else:
        ap.error("one of --synthetic / --npz")
    cam.rows, cam.cols = int(frames[0].shape[0]), int(frames[0].shape[1])
    out = sys.stdout if a.out == "-" else open(a.out, "w")
    res = replay(frames, ts, its, gyro, acc, cam, det, mapc, core, device=a.device, out=out)
    if out is not sys.stdout:
        out.close()
    print(f"replayed {len(frames)} frames, {len(res)} odometry lines", file=sys.stderr)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
