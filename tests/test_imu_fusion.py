"""the IMU fusion block on the device (rbv_fusion.cu) against the UNMODIFIED reference (oracle/_ref run
through its deterministic pipeline with synthetic IMU samples, tests/imu_ref_run.py).

Parity definition. The block is float32 arithmetic in the reference's statement order, but two pieces cannot be bit-exact:
sinf/cosf/tanf/atan2f are CUDA's instead of glibc's, and TooN::SVD<7> (LAPACK dgesvd_, un-vendored) is an FP64 LDL^T /
Jacobi solve here. The Gauss-Newton filter is iterated to a float32 fixed point, so the difference stays at rounding
level of the quantities involved; the tolerances below are relative to the magnitude of each quantity (norm over the
vector or matrix), with the filter state carried by the device over the whole sequence (no per-frame resynchronisation)."""
from __future__ import annotations

import numpy as np
import pytest

import imu_ref_run
from rebvio_b200 import capi, host, pipeline, synth

pytestmark = pytest.mark.gpu

SEQ, NF = "C5:320x240", 40
# relative tolerances (||a - b|| / max(||b||, floor))
# Measured on B200 over the 39 estimated frames (24 of them with the accelerometer filter active): Rt, R, R0, Rgva, Xgv,
# Xgva, V, P_V, Bg, Av, As come out bit-identical to the reference (0.0); X 1.5e-7, K 2.7e-6, orientation 1.8e-6,
# position 1.4e-6. The bounds leave one order of magnitude for other libm / LAPACK builds of the reference.
TOL = dict(Rt=1e-6, R=1e-6, R0=1e-6, Rgva=1e-6, Xgv=1e-6, Xgva=1e-6, V=1e-6, P_V=1e-6, Bg=1e-6, Av=1e-6, As=1e-6, X=2e-6,
           K=3e-5, orientation=2e-5, position=2e-5)
FLOORS = dict(Av=1e-3, orientation=1e-3, position=1e-3, Bg=1e-4)

_REF = {}


def _ref():
    if "recs" not in _REF:
        _REF["recs"] = imu_ref_run.run_in_subprocess(SEQ, NF)
    return _REF["recs"]


def _rel(a, b, floor=1e-12):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), floor))


def _imu_in(r):
    a = np.zeros(1, capi.IMU_IN)
    a["R"], a["gyro"], a["cacc"], a["dt_s"], a["frame_dt"] = r["imu_R"], r["imu_gyro"], r["imu_cacc"], r["imu_dt_s"], r["frame_dt"]
    return a


def _ctx(lanes=2, mode=capi.REDUCE_ORDERED):
    spec = synth.config(SEQ, n_frames=NF)
    cam, det, mapc, core = capi.defaults()
    cam.fm, cam.cx, cam.cy, cam.rows, cam.cols = spec.cam.fm, spec.cam.cx, spec.cam.cy, spec.cam.rows, spec.cam.cols
    det.keylines_ref, det.keylines_max = spec.keylines_ref, spec.keylines_max
    return host.Context(cam, det, mapc, core, lanes=lanes, reduce_mode=mode), det


def test_fusion_block_follows_reference_teacher_forced():
    """Every estimated frame: the reference's tracker results (Vg, W_Xv, Xv) and integrated IMU go into rbv_imu_prerotate /
    rbv_imu_fuse; all outputs of the block are compared. The filter state lives on the device for the whole sequence."""
    recs = _ref()
    ctx, _ = _ctx()
    ctx.imu_reset()
    worst = {k: 0.0 for k in TOL}
    n_acc = 0
    for r in recs:
        if not r["estimated"]:
            continue
        Rt = ctx.imu_prerotate(_imu_in(r))
        o = ctx.imu_fuse(r["Vg"], r["W_Xv"], r["Xv"])[1]
        got = dict(Rt=Rt[1], R=o["R"], R0=o["R0"], Rgva=o["Rgva"], Xgv=o["Xgv"], Xgva=o["Xgva"], V=o["V"], P_V=o["P_V"],
                   Bg=o["Bg"], Av=o["Av"], As=o["As"], X=o["X"], K=o["K"], orientation=o["orientation"], position=o["position"])
        want = dict(Rt=r["Rt1"], R=r["R"], R0=r["R0"], Rgva=r["Rgva"], Xgv=r["Xgv"], Xgva=r["Xgva"], V=r["V"], P_V=r["P_V"],
                    Bg=r["Bg"], Av=r["Av"], As=r["As"], X=r["sabX"], K=r["K"], orientation=r["orientation"],
                    position=r["position"])
        for k in TOL:
            worst[k] = max(worst[k], _rel(got[k], want[k], FLOORS.get(k, 1e-12)))
        assert int(o["initialized"]) == r["initialized"] and int(o["stopped"]) == (1 if r["stopped"] == 1 else 0)
        n_acc += r["K"] != 1.0
    ctx.close()
    print("worst relative differences:", {k: "%.2e" % v for k, v in worst.items()})
    assert n_acc >= 20          # the accelerometer filter ran on most of the sequence
    for k in TOL:
        assert worst[k] <= TOL[k], (k, worst[k])


def test_track_imu_equals_stage_calls_and_follows_reference():
    """Free running: rbv_track_imu (kernels only, one synchronisation per frame) against the same frame done with one C-ABI
    call per reference member function plus the two fusion stage calls - bit-identical - and against the reference's
    own free-running pipeline (keyline and match counts equal, pose and scale within 0.1 % after 39 frames; measured 2e-4)."""
    recs = _ref()
    spec = synth.config(SEQ, n_frames=NF)
    seq = synth.generate(spec)
    frames = [synth.to_float3(f) for f in seq.frames]
    B = 2
    res = []
    for fused in (True, False):
        ctx, det = _ctx(B)
        if fused:
            ctx.close()
            pipe = pipeline.BatchedPipeline(ctx.cam, det, ctx.mapc, ctx.core, lanes=B, imu_cfg=True)
            ctx = pipe.ctx
        else:
            ctx.imu_reset()
        thr, count = np.full(B, det.threshold, np.float32), np.zeros(B, np.int32)
        old, out_seq, last = None, [], None
        for k in range(NF):
            img = np.ascontiguousarray(np.broadcast_to(frames[k], (B,) + frames[k].shape))
            a = _imu_in(recs[k])
            if fused:
                t = pipe.push(img, ts_us=int(seq.ts_us[k]), imu_in=a)
                new, n = pipe.newest_map(), t["keylines"]
                if k:
                    o, nm = t["imu"], t["matches"]
            else:
                if det.gain > 0:   # EdgeDetector::detect threshold servo, edge_detector.cpp:33-36
                    thr = np.clip(thr - np.float32(det.gain) * (np.int32(det.keylines_ref) - count).astype(np.float32),
                                  np.float32(det.min_threshold), np.float32(det.max_threshold)).astype(np.float32)
                new, n, _ = ctx.detect(img, thr, ts_us=int(seq.ts_us[k]))
                count = n
                if old is not None:
                    Rt = ctx.imu_prerotate(a)
                    ctx.buildDistanceField(new)
                    ctx.rotateKeylines(old, Rt)
                    _, Vg, _ = ctx.minimizeVel(old, np.zeros((B, 3), np.float32))
                    ctx.forwardMatch(old, new)
                    ok, Wx, X = ctx.extRotVel(new, Vg)
                    o = ctx.imu_fuse(Vg, Wx, X)
                    ctx.rotateKeylines(old, o["R0"].reshape(B, 3, 3))
                    nm, _ = ctx.directedMatch(new, old, o["V"], o["P_V"].reshape(B, 3, 3), o["Rgva"].reshape(B, 3, 3), 40.0)
                    keep = nm >= 500   # rebvio.cpp:247-252; all lanes stay above the floor in this sequence
                    assert keep.all()
                    ctx.regularize1Iter(new)
                    ctx.updateInverseDepth(o["V"])
                    old.release()
                old = new
            if k:
                out_seq.append((o.tobytes(), np.asarray(nm).tobytes(), np.asarray(n).tobytes(), new.keylines(1)[0].tobytes()))
                last = (o[1].copy(), int(nm[1]), int(n[1]))
        res.append((out_seq, last))
        (pipe if fused else ctx).close()
    for k, (a, b) in enumerate(zip(res[0][0], res[1][0])):
        assert a == b, ("fused and staged frames differ", k + 1)
    o, nm, n = res[0][1]
    r = recs[-1]
    print("free run, frame", NF - 1, "pos", o["position"], r["position"], "ori", o["orientation"], r["orientation"], "K", o["K"], r["K"],
          "matches", nm, r["klm_num"], "keylines", n, r["n_new"])
    assert n == r["n_new"]                       # the front end does not depend on the tracker
    assert abs(nm - r["klm_num"]) <= 0.002 * r["klm_num"]
    assert _rel(o["position"], r["position"]) < 1e-3 and _rel(o["orientation"], r["orientation"], 1e-3) < 1e-3
    assert abs(float(o["K"]) - r["K"]) < 1e-3 * r["K"]


def test_replay_odometry_follows_reference(tmp_path):
    """the headless replay (files in, odometry log out; host IMU integrator + rbv_detect_u8 + rbv_track_imu) on
    the same camera frames and raw IMU samples as the reference's deterministic pipeline: one log line per estimated frame in
    the reference's format, every pose within 1e-4 absolute (metres / radians) of the reference's, free running."""
    import io
    from rebvio_b200 import replay
    recs = _ref()
    spec = synth.config(SEQ, n_frames=NF)
    seq = synth.generate(spec)
    ctx, det = _ctx(1)
    cam, mapc, core = ctx.cam, ctx.mapc, ctx.core
    ctx.close()
    out = io.StringIO()
    res = replay.replay(seq.frames, seq.ts_us, seq.imu_ts_us, seq.gyro, seq.acc, cam, det, mapc, core, out=out)
    lines = out.getvalue().strip().split("\n")
    assert len(res) == NF - 1 == len(lines)
    worst = 0.0
    for (ts, ori, pos, r, nm), line, ref in zip(res, lines, recs[1:]):
        assert ts == ref["ts_us"] and ref["estimated"]
        f = line.split(" ")
        assert len(f) == 7 and int(f[0]) == ts and all(len(x.split(".")[1]) == 6 for x in f[1:])
        worst = max(worst, float(np.abs(ori - ref["orientation"]).max()), float(np.abs(pos - ref["position"]).max()))
        assert abs(nm - ref["klm_num"]) <= max(2, 0.002 * ref["klm_num"])
    print("replay: worst absolute pose difference over %d frames: %.2e" % (len(res), worst))
    assert worst < 1e-4
    assert abs(float(res[-1][3]["K"]) - recs[-1]["K"]) < 1e-3 * recs[-1]["K"]
