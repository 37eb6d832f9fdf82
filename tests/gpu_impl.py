"""Adapter giving the CUDA library (through the C ABI) the method set of oracle/pyref.RefLib so tests/harness.py can
drive it. Inputs are replicated over `lanes` lanes; outputs are read from `lane` and every other lane is asserted
to agree bit for bit (exercises the lane indexing of every kernel)."""
from __future__ import annotations

import numpy as np

from rebvio_b200 import capi, host


class GpuImpl:
    def __init__(self, cam, det, mapc, core, lanes=1, lane=0, reduce_mode=capi.REDUCE_ORDERED, device=0, tune=None):
        conv = lambda cls, src: cls(*[getattr(src, f[0]) for f in cls._fields_])
        self.ctx = host.Context(conv(capi.Camera, cam), conv(capi.DetectorCfg, det), conv(capi.MapCfg, mapc),
                                conv(capi.CoreCfg, core), lanes=lanes, device=device, reduce_mode=reduce_mode)
        for key, value in (tune or {}).items():
            self.ctx.set_tuning(key, value)
        self.lanes, self.lane = lanes, lane
        self.threshold = float(np.float32(det.threshold))
        self.last_count = 0
        self.last_probes = 0
        self._last_detect = None
        self.iterations = int(core.iterations)

    def close(self): self.ctx.close()

    def _same(self, arr, what):
        arr = np.asarray(arr)
        for b in range(self.lanes):
            assert np.ascontiguousarray(arr[b]).tobytes() == np.ascontiguousarray(arr[self.lane]).tobytes(), f"lane {b} differs in {what}"
        return arr[self.lane]

    # maps
    def map_new(self, ts_us=0): return self.ctx.new_map()
    def map_free(self, m): m.release()
    def map_size(self, m): return int(self._same(m.size(), "size"))
    def map_threshold(self, m): return float(self._same(m.threshold(), "threshold"))
    def map_set_threshold(self, m, t): m.set_threshold(t)
    def map_get(self, m): return m.keylines(self.lane)[0]
    def map_pixels(self, m): return m.keylines(self.lane)[1]

    def map_set(self, m, kl, pixel_idx=None):
        for b in range(self.lanes):
            m.upload(b, kl, pixel_idx)

    # front end
    def get_threshold(self): return self.threshold
    def set_threshold(self, t): self.threshold = float(np.float32(t))
    def set_last_count(self, n): self.last_count = n

    def detect(self, img, ts_us=0):
        self.threshold = self.ctx.servo_threshold(self.threshold, self.last_count)
        imgs = np.broadcast_to(np.asarray(img, np.float32), (self.lanes,) + img.shape)
        m, n, at = self.ctx.detect(imgs, self.threshold, ts_us)
        self.last_count = int(self._same(n, "detect count"))
        self._same(at, "auto threshold")
        self._last_detect = m
        return m

    def detector_mask(self): return self._last_detect.mask(self.lane)
    def scale_space_images(self): return [self.ctx.read_image(w, self.lane) for w in range(4)]

    # tracker
    def build_distance_field(self, m): self.ctx.buildDistanceField(m)
    def field(self): return self.ctx.read_field(self.lane)
    def rotate(self, m, R): self.ctx.rotateKeylines(m, np.asarray(R, np.float32).reshape(9))
    def quantile(self, m, pct=0.9, bins=100): return float(self._same(self.ctx.estimateQuantile(m, pct, bins), "quantile"))

    def minimize_vel_trace(self, m, vel, iterations=None, want_residuals=True):
        sc, v, Rv = self.ctx.minimizeVel(m, vel)
        self._same(v, "Vg"); self._same(Rv, "P_Vg")
        n = self.map_size(m)
        return (float(self._same(sc, "score")), v[self.lane].copy(), Rv[self.lane].copy(), self.ctx.lm_trace(self.lane),
                m.residuals(self.lane)[:n])

    def minimize_vel(self, m, vel):
        s, v, Rv, _, _ = self.minimize_vel_trace(m, vel)
        return s, v, Rv

    def try_vel(self, m, vel, sigma_rho_min, residuals=None):
        sc, JtJ, JtF = self.ctx.tryVel(m, vel, sigma_rho_min)
        return float(self._same(sc, "score")), self._same(JtJ, "JtJ").copy(), self._same(JtF, "JtF").copy()

    def forward_match(self, old, new): return int(self._same(self.ctx.forwardMatch(old, new), "fwd count"))

    def ext_rot_vel(self, new, vel):
        ok, Wx, X = self.ctx.extRotVel(new, vel)
        return int(self._same(ok, "ok")), self._same(Wx, "Wx").copy(), self._same(X, "X").copy()

    def directed_match(self, new, old, vel, Rvel, Rback, max_radius=40.0):
        n, kf = self.ctx.directedMatch(new, old, vel, np.asarray(Rvel, np.float32).reshape(9),
                                       np.asarray(Rback, np.float32).reshape(9), max_radius)
        self.last_probes = self.ctx.probe_count(self.lane)
        return int(self._same(n, "directed count")), int(self._same(kf, "kf count"))

    def regularize(self, m): return int(self._same(self.ctx.regularize1Iter(m), "reg count"))
    def update_inverse_depth(self, vel): self.ctx.updateInverseDepth(vel)
