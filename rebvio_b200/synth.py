"""Seeded synthetic edge-rich mono sequences + analytic IMU.

No datasets are available offline, so every test and benchmark input is rendered here on the CPU:
a procedural texture (random-orientation step edges + checker + band-limited noise) is mapped on one or
more planes and sampled through a pinhole camera by inverse ray casting with bilinear interpolation.
Frames are 8-bit mono; `to_float3` applies the `u8 -> float32 * 3.0` conversion of Rebvio::imageCallback
(reference rebvio/src/rebvio.cpp:43) so the hot path sees values in [0, 765], already undistorted
(pinhole, no lens distortion -> cv::undistort at rebvio.cpp:45 is the identity for these frames).

The IMU samples follow the conventions of IntegratedImu::add (reference rebvio/include/rebvio/types/imu.hpp:
51-75): gyro/acc are expressed in the IMU frame, gyro_imu = R_c2i * w_cam, acc_imu = R_c2i * f_cam, with the
EuRoC cam0 extrinsics hard-coded in rebvio/include/rebvio/camera.hpp:39-40.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# EuRoC cam0 intrinsics / extrinsics as hard-coded in the reference Camera (camera.hpp:26-40)
EUROC_FX, EUROC_FY = 458.654, 457.296
EUROC_FM = np.float32(0.5) * (np.float32(EUROC_FX) + np.float32(EUROC_FY))
EUROC_CX, EUROC_CY = 367.215, 248.375
R_C2I = np.array([[0.0148655429818, -0.999880929698, 0.00414029679422],
                  [0.999557249008, 0.0149672133247, 0.025715529948],
                  [-0.0257744366974, 0.00375618835797, 0.999660727178]], dtype=np.float64)
T_C2I = np.array([-0.0216401454975, -0.064676986768, 0.00981073058949], dtype=np.float64)


@dataclass
class CameraSpec:
    fm: float
    cx: float
    cy: float
    rows: int
    cols: int


@dataclass
class SequenceSpec:
    """One of the named configurations C1..C5 of SURVEY.md §8(d)."""
    name: str
    cam: CameraSpec
    n_frames: int
    seed: int
    planes: str = "single"        # "single" fronto-parallel plane | "room" = floor + two walls
    n_edges: int = 64
    checker: tuple = (37, 29)
    fps: float = 20.0
    imu_hz: float = 200.0
    keylines_ref: int = 12000
    keylines_max: int = 16000
    extra: dict = field(default_factory=dict)


def config(name: str, n_frames: int | None = None, seed: int | None = None) -> SequenceSpec:
    """Named workloads. C4 (batched) uses the C2 generator with seeds 1000 + 64*g + i."""
    if name == "C1":  # 640x480 textured plane, reference CPU-runnable case
        s = SequenceSpec("C1", CameraSpec(450.0, 320.0, 240.0, 480, 640), 200, 1)
    elif name == "C2":  # EuRoC-shaped 752x480, single stream latency
        s = SequenceSpec("C2", CameraSpec(float(EUROC_FM), EUROC_CX, EUROC_CY, 480, 752), 200, 2, planes="room")
    elif name == "C3":  # 1280x720 dense-edge scene, ~20k keylines
        s = SequenceSpec("C3", CameraSpec(900.0, 640.0, 360.0, 720, 1280), 100, 3, n_edges=256, checker=(11, 9),
                         keylines_ref=20000, keylines_max=32000)
    elif name.startswith("C5:"):  # kernel sweep: "C5:<cols>x<rows>"
        cols, rows = (int(v) for v in name[3:].split("x"))
        kmax = -(-rows * cols // 22)
        s = SequenceSpec(name, CameraSpec(0.6 * cols, cols / 2.0, rows / 2.0, rows, cols), 20, 5,
                         keylines_ref=int(0.75 * kmax), keylines_max=kmax)
    else:
        raise ValueError(f"unknown config {name}")
    if n_frames is not None:
        s.n_frames = n_frames
    if seed is not None:
        s.seed = seed
    return s


def make_texture(seed: int, size: int = 2048, n_edges: int = 64, checker=(37, 29)) -> np.ndarray:
    """Procedural texture, float32 in [0,255]: random step edges (contrast >= 40), checker (+-25), noise."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32)
    tex = np.full((size, size), 128.0, dtype=np.float32)
    for _ in range(n_edges):
        ang = rng.uniform(0, np.pi)
        nx, ny = np.cos(ang), np.sin(ang)
        off = rng.uniform(0.1, 0.9) * size
        contrast = rng.uniform(40, 90) * rng.choice([-1.0, 1.0])
        # a band (two parallel step edges) keeps the mean bounded
        width = rng.uniform(20, 160)
        d = (xx - size / 2) * nx + (yy - size / 2) * ny + size / 2 - off
        tex += np.where((d > 0) & (d < width), np.float32(contrast), np.float32(0))
    cw, ch = checker
    chk = ((np.floor(xx / cw) + np.floor(yy / ch)) % 2).astype(np.float32)
    tex += (chk * 2 - 1) * 25.0
    # band-limited noise: white noise smoothed by a small box filter (sigma ~ 6 after smoothing gain)
    noise = rng.normal(0, 1, (size, size)).astype(np.float32)
    k = 5
    c = np.cumsum(np.cumsum(np.pad(noise, ((k, 0), (k, 0))), 0), 1)
    sm = (c[k:, k:] - c[:-k, k:] - c[k:, :-k] + c[:-k, :-k]) / k
    tex += 6.0 * sm[:size, :size]
    tex = 128.0 + (tex - tex.mean()) * (60.0 / max(tex.std(), 1e-3))
    return np.clip(tex, 0, 255).astype(np.float32)


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


class Trajectory:
    """p(t) = A o sin(2 pi f t + phi); attitude the same form (small angles). World: x right, y down, z fwd."""
    A = np.array([0.05, 0.03, 0.04])
    F = np.array([0.5, 0.7, 0.3])
    ATT = np.deg2rad(np.array([2.0, 1.5, 3.0]))
    FA = np.array([0.4, 0.6, 0.25])

    def __init__(self, seed: int):
        rng = np.random.RandomState(seed + 7919)
        self.phi = rng.uniform(0, 2 * np.pi, 3)
        self.phia = rng.uniform(0, 2 * np.pi, 3)

    def pos(self, t):
        return self.A * np.sin(2 * np.pi * self.F * t + self.phi)

    def acc(self, t):
        w = 2 * np.pi * self.F
        return -self.A * w * w * np.sin(w * t + self.phi)

    def R_wc(self, t):
        a = self.ATT * np.sin(2 * np.pi * self.FA * t + self.phia)
        return _rot(a[0], a[1], a[2])

    def omega_cam(self, t, h=1e-4):
        """body angular velocity in the camera frame: R' = R [w]x"""
        R0, R1 = self.R_wc(t - h), self.R_wc(t + h)
        dR = R0.T @ R1
        w = np.array([dR[2, 1] - dR[1, 2], dR[0, 2] - dR[2, 0], dR[1, 0] - dR[0, 1]]) / 2.0
        return w / (2 * h)


def _planes(kind: str):
    """(normal n, offset d0 with n.X = d0, tex axis u, tex axis v, tex origin shift) per plane."""
    if kind == "single":
        return [(np.array([0.0, 0.0, 1.0]), 2.0, np.array([1.0, 0, 0]), np.array([0, 1.0, 0]), (0.5, 0.5))]
    if kind == "room":  # back wall + floor + left wall
        return [(np.array([0.0, 0.0, 1.0]), 2.5, np.array([1.0, 0, 0]), np.array([0, 1.0, 0]), (0.5, 0.5)),
                (np.array([0.0, 1.0, 0.0]), 0.9, np.array([1.0, 0, 0]), np.array([0, 0, 1.0]), (0.3, 0.2)),
                (np.array([-1.0, 0.0, 0.0]), 1.6, np.array([0, 0, 1.0]), np.array([0, 1.0, 0]), (0.2, 0.6))]
    raise ValueError(kind)


def render(tex: np.ndarray, cam: CameraSpec, R_wc: np.ndarray, p_w: np.ndarray, planes: str, tex_scale: float) -> np.ndarray:
    """Ray-cast the textured planes; returns u8 [rows, cols]."""
    size = tex.shape[0]
    v, u = np.mgrid[0:cam.rows, 0:cam.cols].astype(np.float64)
    d_c = np.stack([(u - cam.cx) / cam.fm, (v - cam.cy) / cam.fm, np.ones_like(u)], -1)
    d_w = d_c @ R_wc.T
    best_s = np.full(u.shape, np.inf)
    tu = np.zeros_like(u)
    tv = np.zeros_like(u)
    for n, d0, ua, va, (ou, ov) in _planes(planes):
        denom = d_w @ n
        with np.errstate(divide="ignore", invalid="ignore"):
            s = (d0 - n @ p_w) / denom
        ok = (denom > 1e-9) & (s > 0) & (s < best_s)
        X = p_w + s[..., None] * d_w
        tu = np.where(ok, X @ ua * tex_scale + ou * size, tu)
        tv = np.where(ok, X @ va * tex_scale + ov * size, tv)
        best_s = np.where(ok, s, best_s)
    tu = np.mod(tu, size - 1)
    tv = np.mod(tv, size - 1)
    x0 = np.floor(tu).astype(np.int64)
    y0 = np.floor(tv).astype(np.int64)
    fx = (tu - x0).astype(np.float32)
    fy = (tv - y0).astype(np.float32)
    a = tex[y0, x0] * (1 - fx) + tex[y0, x0 + 1] * fx
    b = tex[y0 + 1, x0] * (1 - fx) + tex[y0 + 1, x0 + 1] * fx
    img = a * (1 - fy) + b * fy
    img = np.where(np.isfinite(best_s), img, 128.0)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def to_float3(img_u8: np.ndarray) -> np.ndarray:
    """u8 -> float32 * 3.0 (reference rebvio.cpp:43: convertTo(CV_32FC1, 3.0)); values 0..765."""
    return img_u8.astype(np.float32) * np.float32(3.0)


@dataclass
class Sequence:
    spec: SequenceSpec
    frames: list           # list of u8 [rows, cols]
    ts_us: np.ndarray      # uint64 [n_frames]
    imu_ts_us: np.ndarray  # uint64 [n_imu]
    gyro: np.ndarray       # float32 [n_imu, 3], IMU frame
    acc: np.ndarray        # float32 [n_imu, 3], IMU frame
    R_rel: np.ndarray      # float64 [n_frames, 3, 3]: true R_wc(k-1)^T R_wc(k) (identity for k = 0)


def generate(spec: SequenceSpec) -> Sequence:
    rng = np.random.RandomState(spec.seed)
    tex = make_texture(spec.seed, n_edges=spec.n_edges, checker=spec.checker)
    traj = Trajectory(spec.seed)
    tex_scale = spec.cam.fm / 2.0  # ~1 texel per pixel at Z = 2 m
    t0_us = 1_000_000
    dt_us = int(round(1e6 / spec.fps))
    ts = np.array([t0_us + k * dt_us for k in range(spec.n_frames)], dtype=np.uint64)
    frames, R_rel = [], []
    R_prev = None
    for k in range(spec.n_frames):
        t = (int(ts[k]) - t0_us) * 1e-6
        R = traj.R_wc(t)
        frames.append(render(tex, spec.cam, R, traj.pos(t), spec.planes, tex_scale))
        R_rel.append(np.eye(3) if R_prev is None else R_prev.T @ R)
        R_prev = R
    # IMU: imu_hz samples, first one half a period after t0 so that every frame interval holds >= 2 samples
    # (IntegratedImu::get divides by n-1, imu.hpp:81)
    imu_dt_us = int(round(1e6 / spec.imu_hz))
    imu_ts = np.arange(t0_us - 10 * imu_dt_us + imu_dt_us // 2, int(ts[-1]) + imu_dt_us, imu_dt_us, dtype=np.uint64)
    g_w = np.array([0.0, 9.81, 0.0])
    bias_g = np.array([0.0188, 0.0037, 0.0776]) * 1e-2
    gyro, acc = [], []
    for tu in imu_ts:
        t = (int(tu) - t0_us) * 1e-6
        w_c = traj.omega_cam(t)
        f_c = traj.R_wc(t).T @ (traj.acc(t) - g_w)
        gyro.append(R_C2I @ w_c + bias_g + rng.normal(0, 1.7e-4, 3))
        acc.append(R_C2I @ f_c + rng.normal(0, 2e-3, 3))
    return Sequence(spec, frames, ts, imu_ts, np.asarray(gyro, np.float32), np.asarray(acc, np.float32),
                    np.asarray(R_rel))


# ---- batched renderer (torch; CPU or CUDA) for benchmark-sized workloads ------------------------------------------
def render_batch_torch(tex, cam: CameraSpec, R_wc, p_w, planes: str, tex_scale: float, device="cpu", chunk=32):
    """Same ray-cast as render() for a batch of poses: R_wc [M,3,3], p_w [M,3] (numpy float64) -> u8 [M,rows,cols]
    torch tensor on `device`. Arithmetic is float64 like render(); results agree with render() to +-1 grey level
    (tests/test_host_logic.py). Used to pre-stage bench inputs (64 sequences per GPU), never inside a timed region."""
    import torch
    size = tex.shape[0]
    dev = torch.device(device)
    f64 = torch.float64
    t_tex = torch.as_tensor(tex, dtype=torch.float32, device=dev)
    v, u = torch.meshgrid(torch.arange(cam.rows, dtype=f64, device=dev), torch.arange(cam.cols, dtype=f64, device=dev), indexing="ij")
    d_c = torch.stack([(u - cam.cx) / cam.fm, (v - cam.cy) / cam.fm, torch.ones_like(u)], -1)          # [r,c,3]
    R_all = torch.as_tensor(np.asarray(R_wc), dtype=f64, device=dev)
    p_all = torch.as_tensor(np.asarray(p_w), dtype=f64, device=dev)
    out = torch.empty((R_all.shape[0], cam.rows, cam.cols), dtype=torch.uint8, device=dev)
    pl = [(torch.as_tensor(n, dtype=f64, device=dev), float(d0), torch.as_tensor(ua, dtype=f64, device=dev),
           torch.as_tensor(va, dtype=f64, device=dev), ou, ov) for n, d0, ua, va, (ou, ov) in _planes(planes)]
    for s0 in range(0, R_all.shape[0], chunk):
        R, p = R_all[s0:s0 + chunk], p_all[s0:s0 + chunk]
        d_w = torch.einsum("rck,mjk->mrcj", d_c, R)                                                     # d_c @ R^T
        best = torch.full(d_w.shape[:3], float("inf"), dtype=f64, device=dev)
        tu = torch.zeros_like(best)
        tv = torch.zeros_like(best)
        for n, d0, ua, va, ou, ov in pl:
            denom = d_w @ n
            s = (d0 - (p @ n))[:, None, None] / denom
            ok = (denom > 1e-9) & (s > 0) & (s < best)
            X = p[:, None, None, :] + s[..., None] * d_w
            tu = torch.where(ok, (X @ ua) * tex_scale + ou * size, tu)
            tv = torch.where(ok, (X @ va) * tex_scale + ov * size, tv)
            best = torch.where(ok, s, best)
        tu = torch.remainder(tu, size - 1)
        tv = torch.remainder(tv, size - 1)
        x0 = torch.floor(tu).long()
        y0 = torch.floor(tv).long()
        fx = (tu - x0).float()
        fy = (tv - y0).float()
        a = t_tex[y0, x0] * (1 - fx) + t_tex[y0, x0 + 1] * fx
        b = t_tex[y0 + 1, x0] * (1 - fx) + t_tex[y0 + 1, x0 + 1] * fx
        img = a * (1 - fy) + b * fy
        img = torch.where(torch.isfinite(best), img, torch.full_like(img, 128.0))
        out[s0:s0 + chunk] = torch.clamp(torch.round(img), 0, 255).to(torch.uint8)
    return out


def batched_sequences(spec: SequenceSpec, seeds, n_frames: int, device="cpu", n_textures: int = 4):
    """`len(seeds)` independent sequences of `spec`'s shape (the C4 batched-replay workload of SURVEY.md 8(d)):
    trajectory per seed; the procedural texture is one of `n_textures` (seed % n_textures) to keep set-up time bounded.
    -> frames u8 torch [n_frames, L, rows, cols] on `device`, R_rel float32 numpy [n_frames, L, 3, 3] (identity at 0)."""
    import torch
    L = len(seeds)
    tex_scale = spec.cam.fm / 2.0
    dt = 1.0 / spec.fps
    frames = torch.empty((n_frames, L, spec.cam.rows, spec.cam.cols), dtype=torch.uint8, device=device)
    R_rel = np.zeros((n_frames, L, 3, 3), np.float32)
    texs = {}
    for i, seed in enumerate(seeds):
        tkey = seed % n_textures
        if tkey not in texs:
            texs[tkey] = make_texture(9000 + tkey, n_edges=spec.n_edges, checker=spec.checker)
        traj = Trajectory(seed)
        R = np.stack([traj.R_wc(k * dt) for k in range(n_frames)])
        p = np.stack([traj.pos(k * dt) for k in range(n_frames)])
        frames[:, i] = render_batch_torch(texs[tkey], spec.cam, R, p, spec.planes, tex_scale, device=device)
        R_rel[0, i] = np.eye(3)
        for k in range(1, n_frames):
            R_rel[k, i] = R[k - 1].T @ R[k]
    return frames, R_rel
