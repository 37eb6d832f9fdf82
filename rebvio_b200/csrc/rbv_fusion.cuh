// rbv_fusion.cuh — per-lane state of the IMU fusion block (rbv_fusion.cu). One record per lane in device memory; it is
// the union of what the reference keeps in Rebvio members (imu_state_, sab_state_; rebvio.hpp:92-101), in the locals
// of Rebvio::stateEstimationProcess (rebvio.cpp:95-115) and in the function-local statics of
// Core::estimateLs4Acceleration / estimateMeanAcceleration (core.cpp:292-298, 340-343) — per handle and per lane here,
// so that many sequences can share a process.
#pragma once

struct ImuLane {
  float Bg[3];          // gyro bias
  float W_Bg[9];        // its information matrix
  float RGBias[9];      // gyro-bias observation noise of the previous frame (used by the bias initialisation)
  float R[9];           // inter-frame rotation of the current frame (gyro, bias-corrected; written by k_imu_pre)
  float lsV[5][3];      // estimateLs4Acceleration statics V, V0, V1, V2, V3
  float lsDt[4];
  float maA[4][3];      // estimateMeanAcceleration statics A, A0, A1, A2
  float Av[3], As[3];
  float X[7], P[49];    // scale-attitude-bias filter
  float g_est[3], b_est[3];
  float u_est[3], R_global[9], Pos[3];
  float K, P_Kp;
  float gyro_init[3], g_init[3];
  int num_gyro_init, initialized, num_frames, stopped;
};
