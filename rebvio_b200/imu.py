"""Host side of the IMU path: the sample integrator and the odometry record / log line.

The reference integrates the IMU samples of a frame interval on the host, inside the edge map that the frame produced
(types::IntegratedImu, reference rebvio/include/rebvio/types/imu.hpp:33-150; fed by Rebvio::dataAcquisitionProcess,
rebvio.cpp:70-80). That stays host code here as well - a handful of 3x3 products per 5 ms sample - and its result is the
per-frame input record of the device fusion block (capi.IMU_IN -> rbv_track_imu). Arithmetic is float32 in the
reference's statement order.

Odometry (types::Odometry, rebvio.cpp:270-286): orientation = so3 log of the global rotation and the position, both
computed by the fusion kernel; `odometry_line` is the reference's text log format (`ts rx ry rz px py pz`, 6 decimals).
"""
from __future__ import annotations

import numpy as np

from . import capi

f32 = np.float32

# Camera::R_c2i_, t_c2i_ (camera.hpp:41-45: EuRoC cam0 extrinsics)
R_C2I = np.array([[0.0148655429818, -0.999880929698, 0.00414029679422],
                  [0.999557249008, 0.0149672133247, 0.025715529948],
                  [-0.0257744366974, 0.00375618835797, 0.999660727178]], f32)
T_C2I = np.array([-0.0216401454975, -0.064676986768, 0.00981073058949], f32)


def _mv(M, v):
    """float32 matrix x vector with in-order dot products (TooN evaluates them left to right)"""
    out = np.zeros(3, f32)
    for i in range(3):
        s = f32(M[i, 0] * v[0])
        s = f32(s + f32(M[i, 1] * v[1]))
        out[i] = f32(s + f32(M[i, 2] * v[2]))
    return out


def _mm(A, B):
    out = np.zeros((3, 3), f32)
    for i in range(3):
        for j in range(3):
            s = f32(A[i, 0] * B[0, j])
            s = f32(s + f32(A[i, 1] * B[1, j]))
            out[i, j] = f32(s + f32(A[i, 2] * B[2, j]))
    return out


def so3_exp(w):
    """TooN::SO3<float>::exp (so3.h:254-285) with its Taylor branches; double literals promote as in C++"""
    w = np.asarray(w, f32)
    theta_sq = f32(f32(f32(w[0] * w[0]) + f32(w[1] * w[1])) + f32(w[2] * w[2]))
    theta = f32(np.sqrt(theta_sq))
    one_6th, one_20th = f32(1.0 / 6.0), f32(1.0 / 20.0)
    if theta_sq < 1e-8:
        A, B = f32(1.0 - float(f32(one_6th * theta_sq))), f32(0.5)
    elif theta_sq < 1e-6:
        B = f32(0.5 - 0.25 * float(one_6th) * float(theta_sq))
        A = f32(1.0 - float(f32(theta_sq * one_6th)) * (1.0 - float(f32(one_20th * theta_sq))))
    else:
        inv = f32(1.0 / float(theta))
        A = f32(f32(np.sin(theta)) * inv)
        B = f32(f32(f32(1) - f32(np.cos(theta))) * f32(inv * inv))
    R = np.zeros((3, 3), f32)
    wx2, wy2, wz2 = f32(w[0] * w[0]), f32(w[1] * w[1]), f32(w[2] * w[2])
    R[0, 0] = f32(1.0 - float(f32(B * f32(wy2 + wz2))))
    R[1, 1] = f32(1.0 - float(f32(B * f32(wx2 + wz2))))
    R[2, 2] = f32(1.0 - float(f32(B * f32(wx2 + wy2))))
    a, b = f32(A * w[2]), f32(B * f32(w[0] * w[1])); R[0, 1], R[1, 0] = f32(b - a), f32(b + a)
    a, b = f32(A * w[1]), f32(B * f32(w[0] * w[2])); R[0, 2], R[2, 0] = f32(b + a), f32(b - a)
    a, b = f32(A * w[0]), f32(B * f32(w[1] * w[2])); R[1, 2], R[2, 1] = f32(b - a), f32(b + a)
    return R


class IntegratedImu:
    """types::IntegratedImu (imu.hpp:33-150). add() per IMU sample with ts <= frame ts, get() once per frame."""

    def __init__(self, R_c2i=R_C2I, t_c2i=T_C2I):
        self.Rt = np.ascontiguousarray(np.asarray(R_c2i, f32).T)
        self.t = np.asarray(t_c2i, f32)
        self.n, self.init_ts, self.last_ts, self.dt_us = 0, 0, 0, 0
        self.R = np.eye(3, dtype=f32)
        self.gyro = np.zeros(3, f32); self.acc = np.zeros(3, f32)
        self.dgyro = np.zeros(3, f32); self.cacc = np.zeros(3, f32)
        self.gyro_init = np.zeros(3, f32); self.gyro_last = np.zeros(3, f32)

    def add(self, ts_us, gyro, acc):
        """imu.hpp:51-75"""
        gyro, acc, ts_us = np.asarray(gyro, f32), np.asarray(acc, f32), int(ts_us)
        tmp = _mv(self.Rt, gyro)
        if self.last_ts == 0:
            self.n, self.init_ts, self.last_ts = 1, ts_us, ts_us
            dt = f32(0.005)
            self.gyro_init = gyro.copy(); self.gyro_last = gyro.copy()
            self.R = np.eye(3, dtype=f32)
            self.gyro = tmp.copy()
            self.acc = _mv(self.Rt, acc)
            self.dgyro[:] = 0; self.cacc[:] = 0
        else:
            self.n += 1
            dt = f32(float(f32(ts_us - self.last_ts)) / 1000000.0)
            self.gyro = (self.gyro + tmp).astype(f32)
            self.acc = (self.acc + _mv(self.Rt, acc)).astype(f32)
        self.R = _mm(self.R, so3_exp((tmp * dt).astype(f32)))
        self.last_ts = ts_us
        self.gyro_last = gyro.copy()

    def dt_s(self):
        return f32(float(f32(self.dt_us)) / 1000000.0)

    def get(self, frame_dt=0.0):
        """imu.hpp:80-94 -> one capi.IMU_IN record (frame_dt = (new ts - old ts) / 1e6 is the caller's, rebvio.cpp:182)"""
        if self.n < 2:
            raise ValueError("IntegratedImu.get needs at least two samples per frame (imu.hpp:81 divides by n - 1)")
        self.dt_us = (self.last_ts - self.init_ts) // (self.n - 1) * self.n
        self.gyro = (self.gyro / f32(self.n)).astype(f32)
        self.acc = (self.acc / f32(self.n)).astype(f32)
        self.dgyro = (_mv(self.Rt, (self.gyro_last - self.gyro_init).astype(f32)) / self.dt_s()).astype(f32)
        self.cacc = (self.acc + np.cross(self.dgyro, -_mv(self.Rt, self.t)).astype(f32)).astype(f32)
        self.n, self.init_ts, self.last_ts = 0, 0, 0
        rec = np.zeros(1, capi.IMU_IN)
        rec["R"], rec["gyro"], rec["cacc"], rec["dt_s"], rec["frame_dt"] = self.R.reshape(9), self.gyro, self.cacc, self.dt_s(), frame_dt
        return rec


def odometry_line(ts_us, orientation, position):
    """REBVIO_ODOMETRY log line (rebvio.cpp:275-283): `ts rx ry rz px py pz`, 6 decimals"""
    v = [float(x) for x in orientation] + [float(x) for x in position]
    return f"{int(ts_us)} " + " ".join(f"{x:.6f}" for x in v)
