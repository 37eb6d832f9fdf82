"""Batched replay pipeline over the C ABI: B independent sequences ("lanes") advance one frame per `push`.

Per frame it makes exactly the hot-path calls Rebvio's two worker loops make (reference rebvio/src/rebvio.cpp):
  EdgeDetector::detect (:70)                                             -> rbv_detect / rbv_detect_device
  buildDistanceField, rotateKeylines(R^T), minimizeVel, forwardMatch, extRotVel (:142,165,169,172,177) -> rbv_track_g1
  [host] fusion of (Vg, P_Vg, Wx, X) into V, P_V, R0, Rgva               (:183-233; IMU filters are out of scope, see below)
  rotateKeylines(R0), directedMatch, regularize1Iter, updateInverseDepth (:223/232,245,256,259)         -> rbv_track_g2
The IMU/bias filters between the two halves (gyroBiasCorrection, estimateLs4Acceleration, estimateBias; core.cpp:264-414)
are O(1) host math outside the hot path; this replay driver uses the visual-only subset of that block
(rebvio.cpp:193-200 with the gyro prior switched off): R0 = exp(X[3:6]), V = R0 Vg + X[0:3], P_V = inv(Wx)[0:3,0:3],
Rgva^T = R0 R^T. The relative rotation R of each frame is an input (the reference takes it from the gyro integral).
There is no CPU fallback: everything below runs through librebvio_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi, host


def so3_exp_batch(w):
    """Rodrigues formula, w [B,3] float64 -> [B,3,3]"""
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w, axis=1)
    K = np.zeros((len(w), 3, 3))
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0] = -w[:, 2], w[:, 1], w[:, 2]
    K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -w[:, 0], -w[:, 1], w[:, 0]
    small = th < 1e-12
    ths = np.where(small, 1.0, th)
    a = np.where(small, 1.0, np.sin(ths) / ths)
    b = np.where(small, 0.0, (1 - np.cos(ths)) / ths ** 2)
    return np.eye(3)[None] + a[:, None, None] * K + b[:, None, None] * (K @ K)


def visual_fusion(R, Vg, Wx, X):
    """batched host block between the two tracker halves -> float32 (R0 [B,9], V [B,3], P_V [B,9], Rgva [B,9]).
    numpy statement of what rbv_visual_fusion (C ABI, used by rbv_track / BatchedPipeline) computes; kept as its CPU test oracle"""
    B = len(X)
    R0 = so3_exp_batch(np.asarray(X, np.float64)[:, 3:6])
    V = np.einsum("bij,bj->bi", R0, np.asarray(Vg, np.float64)) + np.asarray(X, np.float64)[:, 0:3]
    Wx64 = np.asarray(Wx, np.float64).reshape(B, 6, 6)
    try:
        P = np.linalg.inv(Wx64)[:, :3, :3]
    except np.linalg.LinAlgError:                      # some lane is singular: fall back to lane by lane
        P = np.empty((B, 3, 3))
        for b in range(B):
            try:
                P[b] = np.linalg.inv(Wx64[b])[:3, :3]
            except np.linalg.LinAlgError:
                P[b] = np.eye(3) * 1e10
    Rt = R0 @ np.transpose(np.asarray(R, np.float64).reshape(B, 3, 3), (0, 2, 1))
    Rgva = np.transpose(Rt, (0, 2, 1))
    f = lambda a, n: np.ascontiguousarray(a, np.float32).reshape(B, n)
    return f(R0, 9), f(V, 3), f(P, 9), f(Rgva, 9)


class BatchedPipeline:
    def __init__(self, cam, det, mapc, core, lanes, device=0, reduce_mode=capi.REDUCE_ORDERED, cuda_stream=None, imu_cfg=None,
                 overlap=False):
        """imu_cfg: None = visual-only fusion (rbv_track); True or a capi.ImuCfg = the reference's IMU filters on the device
        (rbv_track_imu; push then takes `imu_in`, the integrated IMU of every lane, instead of R_rel).
        overlap: the reference's two-thread schedule (rebvio.cpp:56-90 next to :92-293) on two CUDA streams: push(frame n+1)
        enqueues the front end of frame n+1 on the context's front-end stream and, next to it, the tracker of the pair
        (n-1, n); the tracker outputs a push returns are therefore those of the PREVIOUS frame (R_rel / imu_in passed to a push
        belong to that pair too, i.e. to the frame pushed one call earlier). Three maps rotate. Results are bit-identical to
        the in-order schedule, one call later.
        overlap=2: the same with the threshold servo on the device (rbv_detect* with threshold = NULL), so that a front end
        needs nothing read back from the previous one: push(frame n+2) enqueues its front end behind the one of frame n+1
        (possibly still running) and tracks the pair (n-1, n); the front-end stream always has work queued next to the
        tracker. Four maps rotate, outputs arrive two calls later, `keylines` are those of the tracked pair's new map."""
        self.ctx = host.Context(cam, det, mapc, core, lanes=lanes, device=device, reduce_mode=reduce_mode)
        self.with_imu = imu_cfg is not None
        if self.with_imu:
            self.ctx.imu_reset(None if imu_cfg is True else imu_cfg)
        if cuda_stream is not None:
            self.ctx.set_stream(cuda_stream)
        self.lanes = lanes
        self.overlap = bool(overlap)
        self.lag = int(overlap)                       # 0 in order, 1 detect next to the previous pair's tracker, 2 front end one more frame ahead
        self.nmaps = 2 + self.lag
        self.maps = [self.ctx.new_map() for _ in range(self.nmaps)]
        self.n_detected = 0
        if self.overlap:
            self.ctx.set_tuning(capi.TUNE_FE_STREAM, 1)
            self.ctx.set_tuning(capi.TUNE_FIELD_IN_FE, 1)   # the new map's distance field follows its detect on the front-end stream
        self.cur = 0            # index of the map the next detect writes
        self.have_old = False
        self.threshold = np.full(lanes, np.float32(det.threshold), np.float32)   # config_->threshold per lane
        self.count = np.zeros(lanes, np.int32)                                   # EdgeDetector::keylines_count_
        self.min_matches = int(core.global_min_matches_threshold)
        self.frames_done = 0
        self._ptrs = (C.c_void_p * lanes)()

    def close(self):
        for m in self.maps:
            m.release()
        self.ctx.close()

    def _servo(self):
        """edge_detector.cpp:33-36 for every lane, in float32 like rbv_servo_threshold"""
        det = self.ctx.det
        if det.gain > 0:
            t = self.threshold - np.float32(det.gain) * (np.int32(det.keylines_ref) - self.count).astype(np.float32)
            self.threshold = np.clip(t, np.float32(det.min_threshold), np.float32(det.max_threshold)).astype(np.float32)

    def _host_ptrs(self, frames):
        base, stride = frames.ctypes.data, frames.strides[0]
        ptrs = (C.c_void_p * self.lanes)()
        for b in range(self.lanes):
            ptrs[b] = base + b * stride
        return ptrs

    def prefetch(self, frames):
        """start the H2D of the NEXT push's frames (uint8 or float32 [B,rows,cols]) on the copy stream; it overlaps the
        tracker of the current frame. Pass the same array to the next push."""
        capi.check(self.ctx.L.rbv_prefetch_frames(self.ctx.h, self._host_ptrs(frames), 0, int(frames.dtype == np.uint8)))

    def push(self, frames=None, R_rel=None, device_ptr=None, ts_us=None, next_frames=None, imu_in=None):
        """One frame per lane. `frames`: host uint8 [B,rows,cols] as the camera delivers them (converted x3 on the device,
        rebvio.cpp:43) or host float32 (0..765; pageable or pinned) — or `device_ptr`: float32 frames already in device
        memory. R_rel [B,3,3]: rotation of the camera since the previous frame. `next_frames`: the frames of the NEXT
        push; their upload is started right after this frame's front end so that it overlaps the tracker.
        Returns a dict of per-lane outputs (keyline counts always; velocities/matches once two frames are in)."""
        if self.overlap:
            return self._push_overlapped(frames, R_rel, device_ptr, ts_us, next_frames, imu_in)
        ctx, B = self.ctx, self.lanes
        self._servo()
        new = self.maps[self.cur]
        # with an old map in hand the detect call takes no outputs (no device synchronisation): the new map's keyline count and
        # auto threshold come back with the tracker's results
        n = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
        pn, pat = (None, None) if self.have_old else (capi.ptr(n), capi.ptr(at))
        ts = np.ascontiguousarray(np.broadcast_to(np.asarray(0 if ts_us is None else ts_us, np.uint64), (B,)))
        if device_ptr is not None:
            capi.check(ctx.L.rbv_detect_device(ctx.h, new.h, C.c_void_p(device_ptr), capi.ptr(ts), capi.ptr(self.threshold), pn, pat))
        else:
            fn = ctx.L.rbv_detect_u8 if frames.dtype == np.uint8 else ctx.L.rbv_detect
            capi.check(fn(ctx.h, new.h, self._host_ptrs(frames), 0, capi.ptr(ts), capi.ptr(self.threshold), pn, pat))
        if next_frames is not None:
            self.prefetch(next_frames)
        out = {"keylines": n, "auto_threshold": at}
        if self.have_old:
            old = self.maps[1 - self.cur]
            if self.with_imu:   # prerotate, g1, IMU fusion kernel, g2: kernels only (imu_in: capi.IMU_IN records per lane)
                out.update(ctx.track_imu(old, new, imu_in, 40.0, self.min_matches))
            else:
                out.update(ctx.track(old, new, R_rel, 40.0, self.min_matches))   # g1, host fusion (rbv_visual_fusion), g2
        self.count = out["keylines"]
        self.have_old = True
        self.cur = 1 - self.cur
        self.frames_done += 1
        return out

    def _push_overlapped(self, frames, R_rel, device_ptr, ts_us, next_frames, imu_in):
        ctx, B, lag, M = self.ctx, self.lanes, self.lag, self.nmaps
        k = self.n_detected
        new = self.maps[k % M]
        if lag == 1:
            self._servo()                               # with the keyline counts of the previous detect (edge_detector.cpp:33-36)
            thr = capi.ptr(self.threshold)
        else:
            thr = None                                  # servo on the device: nothing of the previous detect is read back
        ts = np.ascontiguousarray(np.broadcast_to(np.asarray(0 if ts_us is None else ts_us, np.uint64), (B,)))
        # front end of frame k on the front-end stream: no outputs, no synchronisation
        if device_ptr is not None:
            capi.check(ctx.L.rbv_detect_device(ctx.h, new.h, C.c_void_p(device_ptr), capi.ptr(ts), thr, None, None))
        else:
            fn = ctx.L.rbv_detect_u8 if frames.dtype == np.uint8 else ctx.L.rbv_detect
            capi.check(fn(ctx.h, new.h, self._host_ptrs(frames), 0, capi.ptr(ts), thr, None, None))
        if next_frames is not None:
            self.prefetch(next_frames)
        out = {}
        if k >= lag + 1:                                # tracker of the pair (k-lag-1, k-lag) on the tracker stream, next to the front end(s)
            old, mid = self.maps[(k - lag - 1) % M], self.maps[(k - lag) % M]
            if self.with_imu:
                out.update(ctx.track_imu(old, mid, imu_in, 40.0, self.min_matches))
            else:
                out.update(ctx.track(old, mid, R_rel, 40.0, self.min_matches))
            out["tracked_keylines"] = out["keylines"]
        if lag == 1:
            n = np.zeros(B, np.int32); at = np.zeros(B, np.float32)
            capi.check(ctx.L.rbv_detect_wait(ctx.h, new.h, capi.ptr(n), capi.ptr(at), None))
            out["keylines"], out["auto_threshold"] = n, at
            self.count = n
        self.n_detected += 1
        self.frames_done += 1
        return out

    def detect_results(self, back=0):
        """overlap modes: (keyline counts, auto thresholds, thresholds used) of the detect `back` pushes ago; waits for that detect only"""
        m = self.maps[(self.n_detected - 1 - back) % self.nmaps]
        B = self.lanes
        n = np.zeros(B, np.int32); at = np.zeros(B, np.float32); th = np.zeros(B, np.float32)
        capi.check(self.ctx.L.rbv_detect_wait(self.ctx.h, m.h, capi.ptr(n), capi.ptr(at), capi.ptr(th)))
        return n, at, th

    def flush(self, args=()):
        """overlap modes: run the trackers of the pairs still in flight; `args` = one dict(R_rel=... | imu_in=...) per pair, oldest
        first -> list of their outputs"""
        k, lag, M = self.n_detected, self.lag, self.nmaps
        outs = []
        for i, a in zip(range(lag), args):
            j = k + i                                   # the push that would have tracked this pair
            if j < lag + 1:
                continue
            old, mid = self.maps[(j - lag - 1) % M], self.maps[(j - lag) % M]
            outs.append(self.ctx.track_imu(old, mid, a["imu_in"], 40.0, self.min_matches) if self.with_imu
                        else self.ctx.track(old, mid, a["R_rel"], 40.0, self.min_matches))
        return outs

    def newest_map(self):
        if self.overlap:
            return self.maps[(self.n_detected - 1) % self.nmaps]
        return self.maps[1 - self.cur]
