// types.hpp — host-side value types of the reference that travel with the hot path's class API (rebvio/include/rebvio/types/*.hpp):
// Image, Imu, Odometry, ImuStateConfig and the per-map IMU integrator `IntegratedImu` (types/imu.hpp:33-150) that
// `EdgeMap::imu()` hands out (edge_map.hpp:70). Included by adapter.hpp (class API) and rebvio.hpp (pipeline façade).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "rebvio_b200.h"

namespace rebvio_b200 {
namespace types {

struct Image {                       // rebvio::types::Image (types/image.hpp): time stamp + pixel data as the driver delivers it
  uint64_t ts_us = 0;
  int rows = 0, cols = 0;
  std::vector<uint8_t> data;         // rows x cols, dense
};
// stand-in for the CV_32FC1 cv::Mat inside the reference's types::Image (types/image.hpp:19-22) where OpenCV is absent: just enough
// surface for EdgeDetector::detect(ImageT&) - a non-owning view
struct FloatMat {
  int rows = 0, cols = 0;
  size_t step = 0;                   // row pitch in bytes
  float* data = nullptr;
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(reinterpret_cast<char*>(data) + (size_t)r * step); }
};
struct FloatImage {                  // types::Image as the detector sees it: time stamp + float32 frame (x3, undistorted)
  uint64_t ts_us = 0;
  FloatMat data;
};
struct Imu {                         // rebvio::types::Imu (types/imu.hpp:22-28)
  uint64_t ts = 0;                   // [us]
  float gyro[3] = {0, 0, 0};
  float acc[3] = {0, 0, 0};
};
struct Odometry {                    // rebvio::types::Odometry (types/odometry.hpp)
  uint64_t ts_us = 0;
  float orientation[3] = {0, 0, 0};  // so3 log of the global rotation
  float position[3] = {0, 0, 0};
};
struct ImuStateConfig : rbv_imu_cfg { ImuStateConfig() { rbv_default_imu_cfg(this); } };   // imu.hpp:153-168

// float32 3x3 helpers in TooN's evaluation order (operators.hh: dot products accumulate left to right)
inline void mat_vec(const float M[9], const float v[3], float out[3]) {
  for (int i = 0; i < 3; ++i) { float s = M[i * 3] * v[0]; s = s + M[i * 3 + 1] * v[1]; out[i] = s + M[i * 3 + 2] * v[2]; }
}
inline void mat_mat(const float A[9], const float B[9], float out[9]) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { float s = A[i * 3] * B[j]; s = s + A[i * 3 + 1] * B[3 + j]; out[i * 3 + j] = s + A[i * 3 + 2] * B[6 + j]; }
}
// TooN::SO3<float>::exp + rodrigues_so3_exp (so3.h:205-285): Taylor branches near zero, double literals promote
inline void so3_exp(const float w[3], float R[9]) {
  const float one_6th = (float)(1.0 / 6.0), one_20th = (float)(1.0 / 20.0);
  float theta_sq = w[0] * w[0]; theta_sq = theta_sq + w[1] * w[1]; theta_sq = theta_sq + w[2] * w[2];
  const float theta = std::sqrt(theta_sq);
  float A, B;
  if (theta_sq < 1e-8) { A = (float)(1.0 - (double)(one_6th * theta_sq)); B = 0.5f; }
  else if (theta_sq < 1e-6) {
    B = (float)(0.5 - 0.25 * one_6th * theta_sq);
    A = (float)(1.0 - (double)(theta_sq * one_6th) * (1.0 - (double)(one_20th * theta_sq)));
  } else {
    const float inv_theta = (float)(1.0 / theta);
    A = std::sin(theta) * inv_theta;
    B = (1 - std::cos(theta)) * (inv_theta * inv_theta);
  }
  const float wx2 = w[0] * w[0], wy2 = w[1] * w[1], wz2 = w[2] * w[2];
  R[0] = (float)(1.0 - (double)(B * (wy2 + wz2)));
  R[4] = (float)(1.0 - (double)(B * (wx2 + wz2)));
  R[8] = (float)(1.0 - (double)(B * (wx2 + wy2)));
  float a = A * w[2], b = B * (w[0] * w[1]); R[1] = b - a; R[3] = b + a;
  a = A * w[1]; b = B * (w[0] * w[2]); R[2] = b + a; R[6] = b - a;
  a = A * w[0]; b = B * (w[1] * w[2]); R[5] = b - a; R[7] = b + a;
}

// rebvio::types::IntegratedImu (types/imu.hpp:33-150): add() per sample, get() once per frame interval. float32 in the
// reference's statement order; compile the translation unit without FMA contraction (-ffp-contract=off) for bit parity.
class IntegratedImu {
 public:
  IntegratedImu() { reset(); std::memset(gyro_init_, 0, sizeof gyro_init_); std::memset(gyro_last_, 0, sizeof gyro_last_); dt_ = 0; }
  unsigned samples() const { return n_; }
  // imu.hpp:51-75. Rt = R_c2i transposed, row-major
  void add(const Imu& imu, const float Rt[9]) {
    float tmp[3], acc[3], dt;
    mat_vec(Rt, imu.gyro, tmp);
    mat_vec(Rt, imu.acc, acc);
    if (last_ts_ == 0) {
      n_ = 1; init_ts_ = imu.ts; last_ts_ = init_ts_;
      dt = 0.005f;
      for (int i = 0; i < 3; ++i) { gyro_init_[i] = imu.gyro[i]; gyro_[i] = tmp[i]; acc_[i] = acc[i]; dgyro_[i] = 0; cacc_[i] = 0; }
      identity(R_);
    } else {
      ++n_;
      dt = (float)((double)(float)(imu.ts - last_ts_) / 1000000.0);
      for (int i = 0; i < 3; ++i) { gyro_[i] = gyro_[i] + tmp[i]; acc_[i] = acc_[i] + acc[i]; }
    }
    float w[3] = {tmp[0] * dt, tmp[1] * dt, tmp[2] * dt}, E[9], Rn[9];
    so3_exp(w, E);
    mat_mat(R_, E, Rn);
    std::memcpy(R_, Rn, sizeof R_);
    last_ts_ = imu.ts;
    for (int i = 0; i < 3; ++i) gyro_last_[i] = imu.gyro[i];
  }
  // imu.hpp:80-94 -> the per-frame input record of the device fusion block. Needs n >= 2 (the reference divides by n - 1);
  // returns false (record untouched) otherwise.
  bool get(const float Rt[9], const float t_c2i[3], float frame_dt, rbv_imu_in* rec) {
    if (n_ < 2) return false;
    dt_ = (last_ts_ - init_ts_) / (n_ - 1) * n_;
    const float n = (float)n_, dts = dt_s();
    float d[3], rd[3], rt[3];
    for (int i = 0; i < 3; ++i) { gyro_[i] = gyro_[i] / n; acc_[i] = acc_[i] / n; d[i] = gyro_last_[i] - gyro_init_[i]; }
    mat_vec(Rt, d, rd);
    for (int i = 0; i < 3; ++i) dgyro_[i] = rd[i] / dts;
    mat_vec(Rt, t_c2i, rt);
    const float b[3] = {-rt[0], -rt[1], -rt[2]};
    cacc_[0] = acc_[0] + (dgyro_[1] * b[2] - dgyro_[2] * b[1]);
    cacc_[1] = acc_[1] + (dgyro_[2] * b[0] - dgyro_[0] * b[2]);
    cacc_[2] = acc_[2] + (dgyro_[0] * b[1] - dgyro_[1] * b[0]);
    n_ = 0; init_ts_ = 0; last_ts_ = 0;
    std::memcpy(rec->R, R_, sizeof R_);
    for (int i = 0; i < 3; ++i) { rec->gyro[i] = gyro_[i]; rec->cacc[i] = cacc_[i]; }
    rec->dt_s = dts;
    rec->frame_dt = frame_dt;
    return true;
  }
  float dt_s() const { return (float)((double)(float)dt_ / 1000000.0); }

 private:
  static void identity(float R[9]) { for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0f : 0.0f; }
  void reset() {
    n_ = 0; init_ts_ = last_ts_ = 0; identity(R_);
    for (int i = 0; i < 3; ++i) gyro_[i] = acc_[i] = dgyro_[i] = cacc_[i] = 0;
  }
  unsigned n_;
  uint64_t last_ts_, init_ts_, dt_;
  float R_[9], gyro_[3], gyro_init_[3], gyro_last_[3], acc_[3], dgyro_[3], cacc_[3];
};

// REBVIO_ODOMETRY log line (rebvio.cpp:275-283): `ts rx ry rz px py pz`, 6 decimals
inline std::string odometry_line(const Odometry& o) {
  char buf[256];
  std::snprintf(buf, sizeof buf, "%llu %.6f %.6f %.6f %.6f %.6f %.6f", (unsigned long long)o.ts_us, o.orientation[0], o.orientation[1],
                o.orientation[2], o.position[0], o.position[1], o.position[2]);
  return buf;
}

}  // namespace types
}  // namespace rebvio_b200
