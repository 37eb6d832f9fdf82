// rbv_fusion.cu — the per-frame IMU fusion block of Rebvio::stateEstimationProcess on the device:
// gyro-bias initialisation and pre-rotation (rebvio.cpp:146-165), Core::gyroBiasCorrection (core.cpp:264-287),
// estimateLs4Acceleration / estimateMeanAcceleration (core.cpp:289-352), Core::estimateBias with the scale-attitude-bias
// Gauss-Newton filter (core.cpp:355-414, sab_estimator.cpp:21-165), the rotation / translation composition between
// the two halves of the tracker (rebvio.cpp:183-233) and the incremental pose (rebvio.cpp:255-268).
//
// Why on the device: the block sits between extRotVel and the second rotateKeylines, so on the host it costs a device
// round trip in the middle of every frame and, with B lanes, B serial 20-iteration Gauss-Newton solves. Here ONE WARP
// per lane keeps the lane's filter state in HBM, reads the tracker's results where the tracker kernels left them
// (LMState, lane_out) and writes the inputs of the second half (lane_in), so a frame is kernels only.
//
// Arithmetic: float32 in the reference's statement order (TooN evaluates a product element by element as an in-order
// dot product; the double-typed sub-expressions that TooN's scalar promotion creates are kept double). Every small
// matrix lives in shared memory, products are spread over the 32 lanes (one output element per lane and round), the
// Cholesky factorisations (TooN/Cholesky.h:95-118) run on lane 0 and their back-substitutions one column per lane.
// Not bit-exact against the reference by construction: sinf/cosf/tanf/atan2f are CUDA's, and TooN::SVD<7> (LAPACK
// dgesvd_, un-vendored) is replaced by the FP64 LDL^T / Jacobi pair below; the parity tolerance is stated in
// tests/test_imu_fusion.py.
#include <cuda_runtime.h>
#include <math.h>
#include <string.h>

#include "rbv_internal.cuh"
#include "rbv_fusion.cuh"

namespace {

#define FULL 0xFFFFFFFFu
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ double shfl_dd(double v, int src) { return __shfl_sync(FULL, v, src); }

// C[n x m] = A[n x k] * B[k x m]; element (i,j) of A is A[i*ars + j*acs] (a transpose is a stride swap). In-order dot
// products, one output element per lane and round. Ends with a warp barrier.
__device__ void mm(float* C, const float* A, int ars, int acs, const float* B, int brs, int bcs, int n, int k, int m) {
  for (int e = lane_id(); e < n * m; e += 32) {
    const int i = e / m, j = e - i * m;
    float s = A[i * ars] * B[j * bcs];
    for (int q = 1; q < k; ++q) s += A[i * ars + q * acs] * B[q * brs + j * bcs];
    C[e] = s;
  }
  __syncwarp();
}
template <int N, int K, int M>   // compile-time shape: constant divisions, unrolled dot products
__device__ __forceinline__ void mm(float* C, const float* A, int ars, int acs, const float* B, int brs, int bcs) {
#pragma unroll
  for (int e0 = 0; e0 < N * M; e0 += 32) {
    const int e = e0 + lane_id();
    if (e < N * M) {
      const int i = e / M, j = e - i * M;
      float s = A[i * ars] * B[j * bcs];
#pragma unroll
      for (int q = 1; q < K; ++q) s += A[i * ars + q * acs] * B[q * brs + j * bcs];
      C[e] = s;
    }
  }
  __syncwarp();
}
template <class F>
__device__ __forceinline__ void each(int n, F f) {
  for (int e = lane_id(); e < n; e += 32) f(e);
  __syncwarp();
}
template <class F>
__device__ __forceinline__ void one(F f) {
  if (lane_id() == 0) f();
  __syncwarp();
}

// TooN::Cholesky<N,float>(A).get_inverse() (Cholesky.h:95-118 do_compute, :155-181 matrix backsub of the identity).
// `ch` is an N*N workspace. The factorisation is a dependent chain (lane 0); in the back-substitution every column of
// the identity is independent and sees exactly the arithmetic of TooN's row-vector form (lane c = column c).
__device__ void chol_inverse(float* out, const float* A, float* ch, int N) {
  each(N * N, [&](int e) { ch[e] = A[e]; });
  one([&]() {
    for (int col = 0; col < N; ++col) {
      float inv_diag = 1.0f;
      for (int row = col; row < N; ++row) {
        float val = ch[row * N + col];
        for (int c2 = 0; c2 < col; ++c2) val -= ch[c2 * N + col] * ch[row * N + c2];
        if (row == col) {
          ch[row * N + col] = val;
          if (val == 0.0f) return;   // rank deficient: TooN stops here too (the rest keeps the input values)
          inv_diag = 1.0f / val;
        } else {
          ch[col * N + row] = val;
          ch[row * N + col] = val * inv_diag;
        }
      }
    }
  });
  const int c = lane_id();
  if (c < N) {
    float y[7], r[7];
    for (int i = 0; i < N; ++i) {
      float val = (i == c) ? 1.0f : 0.0f;
      for (int j = 0; j < i; ++j) val -= ch[i * N + j] * y[j];
      y[i] = val;
    }
    for (int i = 0; i < N; ++i) y[i] *= (1.0f / ch[i * N + i]);
    for (int i = N - 1; i >= 0; --i) {
      float val = y[i];
      for (int j = i + 1; j < N; ++j) val -= ch[j * N + i] * r[j];
      r[i] = val;
    }
    for (int i = 0; i < N; ++i) out[i * N + c] = r[i];
  }
  __syncwarp();
}

// types::invert (definitions.hpp:40-54): cofactors over TooN::determinant (LU with partial pivoting, rbv_det3)
__device__ void invert3(float* out, const float* in) {
  float o[9];
  o[0] = in[4] * in[8] - in[5] * in[7]; o[1] = in[2] * in[7] - in[1] * in[8]; o[2] = in[1] * in[5] - in[2] * in[4];
  o[3] = in[5] * in[6] - in[3] * in[8]; o[4] = in[0] * in[8] - in[2] * in[6]; o[5] = in[2] * in[3] - in[0] * in[5];
  o[6] = in[3] * in[7] - in[4] * in[6]; o[7] = in[1] * in[6] - in[0] * in[7]; o[8] = in[0] * in[4] - in[1] * in[3];
  const float det = rbv_det3(in);
  for (int i = 0; i < 9; ++i) out[i] = o[i] / det;
}

// TooN::SO3<float>::exp (so3.h:254-285, rodrigues_so3_exp :219-248); the double literals promote as in the reference
__device__ void so3_exp(float* R, const float* w) {
  const float theta_sq = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
  const float theta = sqrtf(theta_sq);
  const float one_6th = (float)(1.0 / 6.0), one_20th = (float)(1.0 / 20.0);
  float A, B;
  if (theta_sq < 1e-8) {
    A = (float)(1.0 - (double)(one_6th * theta_sq));
    B = 0.5f;
  } else if (theta_sq < 1e-6) {
    B = (float)(0.5 - 0.25 * (double)one_6th * (double)theta_sq);
    A = (float)(1.0 - (double)(theta_sq * one_6th) * (1.0 - (double)(one_20th * theta_sq)));
  } else {
    const float inv_theta = (float)(1.0 / (double)theta);
    A = sinf(theta) * inv_theta;
    B = (1 - cosf(theta)) * (inv_theta * inv_theta);
  }
  const float wx2 = w[0] * w[0], wy2 = w[1] * w[1], wz2 = w[2] * w[2];
  R[0] = (float)(1.0 - (double)(B * (wy2 + wz2)));
  R[4] = (float)(1.0 - (double)(B * (wx2 + wz2)));
  R[8] = (float)(1.0 - (double)(B * (wx2 + wy2)));
  { const float a = A * w[2], b = B * (w[0] * w[1]); R[1] = b - a; R[3] = b + a; }
  { const float a = A * w[1], b = B * (w[0] * w[2]); R[2] = b + a; R[6] = b - a; }
  { const float a = A * w[0], b = B * (w[1] * w[2]); R[5] = b - a; R[7] = b + a; }
}

__device__ __forceinline__ void cross3(float* o, const float* a, const float* b) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ float dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ void unit3(float* o, const float* a) {
  const float n = sqrtf(dot3(a, a));
  o[0] = a[0] / n; o[1] = a[1] / n; o[2] = a[2] / n;
}
__device__ void mat3_mul(float* C, const float* A, const float* B) {   // C = A B (C must not alias)
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}
__device__ void mat3_mul_bt(float* C, const float* A, const float* B) {   // C = A B^T
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3] * B[j * 3] + A[i * 3 + 1] * B[j * 3 + 1] + A[i * 3 + 2] * B[j * 3 + 2];
}
__device__ void mat3_vec(float* o, const float* A, const float* v) {
  for (int i = 0; i < 3; ++i) o[i] = A[i * 3] * v[0] + A[i * 3 + 1] * v[1] + A[i * 3 + 2] * v[2];
}
__device__ void mat3t_vec(float* o, const float* A, const float* v) {   // A^T v
  for (int i = 0; i < 3; ++i) o[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
}

// TooN::SO3(a, b): the rotation that takes direction a to direction b (so3.h:78-98)
__device__ void so3_from_to(float* R, const float* a, const float* b) {
  float n[3];
  cross3(n, a, b);
  if (dot3(n, n) == 0.0f) { for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0f : 0.0f; return; }
  unit3(n, n);
  float R1[9], M[9], ua[3], ub[3], t[3];   // R1.T()[k] = column k of R1
  unit3(ua, a); cross3(t, ua, n);
  for (int i = 0; i < 3; ++i) { R1[i * 3] = ua[i]; R1[i * 3 + 1] = n[i]; R1[i * 3 + 2] = t[i]; }
  unit3(ub, b); cross3(t, ub, n);
  for (int i = 0; i < 3; ++i) { M[i * 3] = ub[i]; M[i * 3 + 1] = n[i]; M[i * 3 + 2] = t[i]; }
  mat3_mul_bt(R, M, R1);
}

// TooN::SO3::ln (so3.h:288-332)
__device__ void so3_ln(float* r, const float* M) {
  const float cos_angle = (float)(((double)(M[0] + M[4] + M[8]) - 1.0) * 0.5);
  r[0] = (M[7] - M[5]) / 2; r[1] = (M[2] - M[6]) / 2; r[2] = (M[3] - M[1]) / 2;
  const float sin_abs = sqrtf(dot3(r, r));
  if (cos_angle > M_SQRT1_2) {
    if (sin_abs > 0) { const float f = asinf(sin_abs) / sin_abs; r[0] *= f; r[1] *= f; r[2] *= f; }
  } else if (cos_angle > -M_SQRT1_2) {
    const float f = acosf(cos_angle) / sin_abs;
    r[0] *= f; r[1] *= f; r[2] *= f;
  } else {
    const float angle = (float)(M_PI - (double)asinf(sin_abs));
    const float d0 = M[0] - cos_angle, d1 = M[4] - cos_angle, d2 = M[8] - cos_angle;
    float r2[3];
    if (d0 * d0 > d1 * d1 && d0 * d0 > d2 * d2) { r2[0] = d0; r2[1] = (M[3] + M[1]) / 2; r2[2] = (M[2] + M[6]) / 2; }
    else if (d1 * d1 > d2 * d2) { r2[0] = (M[3] + M[1]) / 2; r2[1] = d1; r2[2] = (M[7] + M[5]) / 2; }
    else { r2[0] = (M[2] + M[6]) / 2; r2[1] = (M[7] + M[5]) / 2; r2[2] = d2; }
    if (dot3(r2, r) < 0) { r2[0] = -r2[0]; r2[1] = -r2[1]; r2[2] = -r2[2]; }
    unit3(r2, r2);
    r[0] = angle * r2[0]; r[1] = angle * r2[1]; r[2] = angle * r2[2];
  }
}

// ---- TooN::SVD<7>(JtJ).backsub(rhs) (sab_estimator.cpp:31-32; double precision, LAPACK dgesvd_ un-vendored) ----
// Same pair as extRotVel's 6x6 solve (rbv_tracker.cu): FP64 LDL^T when the matrix is provably well inside the
// conditioning rule of SVD::backsub (1/s_i zeroed when s_i * 1e9 <= s_max; s_max <= trace, s_min >= 1/||A^-1||_F), the
// Jacobi pseudo-inverse (solve_jacobi8_warp below) otherwise. S = symmetrised matrix in shared memory (double, N*N), b = right-hand side.
// Returns x[lane] for lane < N. All 32 lanes must call.
template <int N>
__device__ bool solve_ldl_warp(const double* S, const double* b, double& xout) {
  const int lane = lane_id();
  double L[N][N], D[N];
  bool ok = true;
  double tr = 0.0;
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = S[j * N + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
    D[j] = d;
    ok = ok && (d > 0.0);
    tr += S[j * N + j];
    const double inv = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double l = S[i * N + j];
#pragma unroll
      for (int k = 0; k < j; ++k) l -= L[i][k] * L[j][k] * D[k];
      L[i][j] = l * inv;
    }
  }
  const int col = lane < N ? lane : 0;
  double z[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double y = (i == col) ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) y -= L[i][k] * z[k];
    z[i] = y;
  }
#pragma unroll
  for (int i = 0; i < N; ++i) z[i] /= D[i];
#pragma unroll
  for (int i = N - 1; i >= 0; --i)
#pragma unroll
    for (int k = i + 1; k < N; ++k) z[i] -= L[k][i] * z[k];
  double f = 0.0, x = 0.0;
#pragma unroll
  for (int k = 0; k < N; ++k) { f += z[k] * z[k]; x += z[k] * b[k]; }
  if (lane >= N) f = 0.0;
#pragma unroll
  for (int o2 = 4; o2; o2 >>= 1) f += __shfl_xor_sync(FULL, f, o2);
  f = shfl_dd(f, 0);
  ok = ok && (sqrt(f) * tr * 4.0 < 1e9);   // false for NaN / Inf as well
  xout = x;
  return ok;
}

// The eigen path: symmetric Jacobi pseudo-inverse, 1/lambda_i zeroed when lambda_i * 1e9 <= lambda_max. It runs 20 times
// per frame on nearly the same matrix and - the filter's prior information spans 1e-2 .. 1e13 - is the path always taken
// once the accelerometer filter is active. Three differences to the cyclic one-warp Jacobi of extRotVel (rbv_tracker.cu):
//  * round-robin ordering on the matrix padded to 8x8 (row/column 7 is zero: one more zero eigenvalue, dropped by the
//    conditioning rule): the 4 rotations of a round touch disjoint index pairs, every lane computes their 4 (c, s)
//    together (independent FP64 chains), lane r < 8 owns row r, row pairs are exchanged with one shuffle per element;
//  * warm start: V of the previous Gauss-Newton iteration diagonalises the new matrix almost completely, so it is
//    transformed first (V^T A V through shared memory) and the sweeps continue from there (1-2 instead of ~8);
//  * the sweeps stop when the off-diagonal norm is below 1e-13 of the matrix norm (see the loop).
// v[8] (row `lane` of V) persists in the caller; warm = 0 starts from the identity. sA: 64 doubles = the symmetric matrix
// padded to 8x8, sV / sT: 64 doubles of workspace each. Returns x[lane] for lane < 7.
struct RotCS { double c, s; };
__device__ __forceinline__ RotCS jacobi_cs(double app, double aqq, double apq) {
  // t = sgn(theta) / (|theta| + sqrt(theta^2 + 1)), theta = (aqq - app) / (2 apq), written as sgn * |b| / (|d| + hypot(d, b)).
  // The angle only has to be good enough for the sweeps to converge (a float32 t leaves 1e-7 of the annihilated element, the
  // next sweep takes care of it); what must hold to FP64 accuracy is c^2 + s^2 = 1, and c = rsqrt(1 + t^2), s = t c does
  // that for any t. So t comes from float32 arithmetic on (d, b) scaled by a power of two (exact) - FP64 divide and square
  // root are ~30-instruction dependent chains, and there are three of them in the textbook formula.
  RotCS r;
  r.c = 1.0; r.s = 0.0;
  if (apq == 0.0) return r;
  const double d = aqq - app, b = 2.0 * apq;
  const int e = (__double2hiint(fmax(fabs(d), fabs(b))) >> 20) & 0x7ff;
  if (e == 0 || e == 0x7ff) return r;                                   // denormal or not finite: leave the pair alone
  const double sc = __hiloint2double((2046 - e) << 20, 0);              // 2^(1023 - e): the larger magnitude lands in [1, 2)
  const double ds = d * sc, bs = b * sc;
  const float df = (float)ds, bf = (float)bs;
  const float tf = fabsf(bf) / (fabsf(df) + sqrtf(df * df + bf * bf));
  double t = ((d >= 0.0) == (b >= 0.0)) ? (double)tf : -(double)tf;     // sign of theta (theta = 0 counts as +)
  // one Newton step on b t^2 + 2 d t - b = 0 (the small root) in FP64, its divide through a float32 reciprocal refined
  // once: t goes from 1e-7 to ~1e-13 relative, which keeps the sweeps quadratically convergent
  const double den = 2.0 * (bs * t + ds);
  double ri = (double)(1.0f / (float)den);
  ri = ri + ri * (1.0 - den * ri);
  t = t - ((bs * t + 2.0 * ds) * t - bs) * ri;
  r.c = rsqrt(t * t + 1.0);
  r.s = t * r.c;
  return r;
}
template <int R>
__device__ __forceinline__ void jacobi_round(double (&a)[8], double (&v)[8], int lane) {
  // round R of the 8-player round robin: (7, R) and ((R+i) % 7, (R-i+7) % 7), i = 1..3
  constexpr int P[4] = {R, (R + 1) % 7, (R + 2) % 7, (R + 3) % 7};
  constexpr int Q[4] = {7, (R + 6) % 7, (R + 5) % 7, (R + 4) % 7};
  RotCS cs[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = P[i] < Q[i] ? P[i] : Q[i], q = P[i] < Q[i] ? Q[i] : P[i];
    cs[i] = jacobi_cs(shfl_dd(a[p], p), shfl_dd(a[q], q), shfl_dd(a[q], p));
  }
  int partner = lane;
  double cm = 1.0, sm = 0.0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = P[i] < Q[i] ? P[i] : Q[i], q = P[i] < Q[i] ? Q[i] : P[i];
    { const double akp = a[p], akq = a[q]; a[p] = cs[i].c * akp - cs[i].s * akq; a[q] = cs[i].s * akp + cs[i].c * akq; }
    { const double vkp = v[p], vkq = v[q]; v[p] = cs[i].c * vkp - cs[i].s * vkq; v[q] = cs[i].s * vkp + cs[i].c * vkq; }
    if (lane == p) { partner = q; cm = cs[i].c; sm = -cs[i].s; }
    if (lane == q) { partner = p; cm = cs[i].c; sm = cs[i].s; }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const double other = shfl_dd(a[k], partner);
    a[k] = cm * a[k] + sm * other;        // row p: c*row_p - s*row_q; row q: s*row_p + c*row_q
  }
}
__device__ double solve_jacobi8_warp(const double* sA, const double* b7, double (&v)[8], bool warm, double* sV, double* sT) {
  const int lane = lane_id();
  const bool act = lane < 8;
  double a[8];
  if (warm) {
    if (act) for (int j = 0; j < 8; ++j) sV[lane * 8 + j] = v[j];
    __syncwarp();
    for (int e = lane; e < 64; e += 32) {
      const int i = e >> 3, j = e & 7;
      double t = 0.0;
      for (int k = 0; k < 8; ++k) t += sA[i * 8 + k] * sV[k * 8 + j];
      sT[e] = t;
    }
    __syncwarp();
    double m[2];
    for (int r = 0; r < 2; ++r) {
      const int e = lane + 32 * r, i = e >> 3, j = e & 7;
      double t = 0.0;
      for (int k = 0; k < 8; ++k) t += sV[k * 8 + i] * sT[k * 8 + j];
      m[r] = t;
    }
    __syncwarp();
    sV[lane] = m[0]; sV[lane + 32] = m[1];          // sV now holds V^T A V
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = act ? 0.5 * (sV[lane * 8 + j] + sV[j * 8 + lane]) : 0.0;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] = act ? sA[lane * 8 + j] : 0.0; v[j] = (act && lane == j) ? 1.0 : 0.0; }
  }
  __syncwarp();
  for (int sweep = 0; sweep < 30; ++sweep) {
    // converged when the off-diagonal norm is below 1e-13 of the matrix norm (1e-26 on the squared masses). The matrix is
    // float32 data (J^T J of the reference, relative accuracy 6e-8) and changes by float32 ulps from one Gauss-Newton
    // iteration to the next, so after the warm start ONE sweep (quadratic: 1e-7 -> 1e-14) lands below the bound, and an
    // unchanged matrix needs none. Eigenvalues that survive the conditioning rule (>= 1e-9 of the largest) are accurate to
    // 1e-4 relative in the worst case and ~1e-7 typically - the accuracy of the data; the fixed point the 20 iterations
    // converge to does not depend on the step accuracy.
    double off = 0.0, dia = 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j > lane) off += a[j] * a[j];
      if (j == lane) dia += a[j] * a[j];
    }
#pragma unroll
    for (int o2 = 4; o2; o2 >>= 1) { off += __shfl_xor_sync(FULL, off, o2); dia += __shfl_xor_sync(FULL, dia, o2); }
    off = shfl_dd(off, 0); dia = shfl_dd(dia, 0);
    if (!(off > 1e-26 * dia)) break;                // also leaves on NaN
    jacobi_round<0>(a, v, lane); jacobi_round<1>(a, v, lane); jacobi_round<2>(a, v, lane); jacobi_round<3>(a, v, lane);
    jacobi_round<4>(a, v, lane); jacobi_round<5>(a, v, lane); jacobi_round<6>(a, v, lane);
  }
  double d = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) if (j == lane) d = a[j];
  double smax = act ? fabs(d) : 0.0;
#pragma unroll
  for (int o2 = 4; o2; o2 >>= 1) smax = fmax(smax, __shfl_xor_sync(FULL, smax, o2));
  smax = shfl_dd(smax, 0);
  // x = sum_i v_i (v_i . b) / lambda_i over the kept eigenpairs (SVD::backsub's conditioning rule, SVD.h:266-274), through
  // shared memory: lane k stores row k of V scaled by b[k], lane i sums column i = v_i . b (k ascending)
  const double bk = lane < 7 ? b7[lane] : 0.0;
  if (act) for (int i = 0; i < 8; ++i) sT[lane * 8 + i] = v[i] * bk;
  __syncwarp();
  if (act) {
    double proj = sT[lane];
    for (int k = 1; k < 8; ++k) proj += sT[k * 8 + lane];
    sV[lane] = (fabs(d) * 1e9 > smax) ? proj / d : 0.0;
  }
  __syncwarp();
  double x = 0.0;
  if (act) for (int i = 0; i < 8; ++i) x += v[i] * sV[i];
  __syncwarp();
  return x;
}

// ---- rebvio.cpp:146-165: bias initialisation and the bias-corrected gyro rotation; one thread per lane ----
__global__ void k_imu_pre(ImuLane* __restrict__ st, const rbv_imu_in* __restrict__ in, rbv_imu_cfg cfg,
                          float* __restrict__ lane_in, rbv_imu_out* __restrict__ out, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  ImuLane& S = st[b];
  const rbv_imu_in I = in[b];
  if (!S.initialized && S.num_frames > 0) {
    if (cfg.init_bias > 0) {
      for (int i = 0; i < 3; ++i) { S.gyro_init[i] += I.gyro[i] * I.dt_s; S.g_init[i] -= I.cacc[i]; }
      if (++S.num_gyro_init > cfg.init_bias_frame_num) {
        const float n = (float)S.num_gyro_init;
        float scaled[9];
        for (int i = 0; i < 9; ++i) scaled[i] = (float)((double)S.RGBias[i] * 1e2);
        for (int i = 0; i < 3; ++i) { S.Bg[i] = S.gyro_init[i] / n; S.X[1 + i] = S.g_init[i] / n; }
        invert3(S.W_Bg, scaled);
        S.initialized = 1;
      }
    } else {
      S.initialized = 1;
      for (int i = 0; i < 3; ++i) S.Bg[i] = cfg.init_bias_guess[i] * I.dt_s;
    }
  }
  float E[9], Rt[9];
  so3_exp(E, S.Bg);
  mat3_mul_bt(Rt, E, I.R);                       // R.T() = SO3(Bg) * R.T()
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) S.R[i * 3 + j] = Rt[j * 3 + i];
  float* li = lane_in + (size_t)b * IN_STRIDE;
  for (int i = 0; i < 9; ++i) { li[IN_R + i] = Rt[i]; out[b].Rt[i] = Rt[i]; }
  for (int i = 0; i < 3; ++i) li[IN_VEL + i] = 0.0f;   // imu_state_.Vg = Zeros, rebvio.cpp:167
}

struct FusionSh {
  float Wx[36], X[6], Vg[3], Xgva[6];
  float R[9], Rgva[9], R0[9];
  float Wg[9], Wb[9], iWgWb[9];
  float Wxb[36], inv6[36], ch[49], X1[6];
  float V[3], P_V[9], P_W[9], Rv[9];
  // estimateBias
  float Xs[7], xp[7], F7[49], Q[49], T7[49], T7b[49], Pp[49], W7[49];
  // SABEstimator::problem
  float F[11], dFda[11], dFdx1[66], Wm[121];
  float M6a[66], M6b[66], JtJ[49], JtF[7];
  double S7[49], rhs[7];
  double A8[64], V8[64], T8[64];
  float as[3], av[3];
  float Pz[9], dP3[9], W3[9], T3[9], dW3[9];
  float G, Rg, K;
  int stopped;
};

// SABEstimator::problem (sab_estimator.cpp:44-163): JtJ (7x7) and JtF (7) at the state s.Xs. s.W7 = inverse of the prior
// covariance s.Pp (constant over the Gauss-Newton iterations; the reference recomputes the same values every time).
// The 11x11 matrices of the reference are block diagonal: W = diag(W3, 1/Rg, W7), P = diag(Pz, Rg, Pp), and dPda / dWda are
// non-zero in the leading 3x3 block only. A dot product that runs over the structural zeros as well adds exact zeros to
// the same partial sums, so the blocks are multiplied directly (same terms, same order, same roundings); only the
// products with the dense 11x6 Jacobian stay dense. Constant entries (zeros of dFdx1, Wm outside the 3x3 block) are
// written once by the caller.
__device__ void sab_problem(FusionSh& s, const float* Rs) {
  one([&]() {
    const float a = s.Xs[0];
    const float* g = s.Xs + 1;
    const float* b = s.Xs + 4;
    const float sa = sinf(a), ca = cosf(a);
    for (int i = 0; i < 11; ++i) { s.F[i] = 0.0f; s.dFda[i] = 0.0f; }
    for (int i = 0; i < 3; ++i) {
      s.F[i] = (s.as[i] + g[i]) * ca - s.av[i] * sa;
      s.dFda[i] = -(s.as[i] + g[i]) * sa - s.av[i] * ca;
    }
    s.dFda[4] = 1.0f;
    s.F[3] = dot3(g, g) - s.G * s.G;
    float f4 = s.Xs[0] - s.xp[0];
    if (f4 > M_PI) f4 = (float)((double)f4 - 2.0 * M_PI);
    else if (f4 < -M_PI) f4 = (float)((double)f4 + 2.0 * M_PI);
    s.F[4] = f4;
    float Rb[9], Rbg[3];
    so3_exp(Rb, b);
    mat3_vec(Rbg, Rb, g);
    for (int i = 0; i < 3; ++i) { s.F[5 + i] = Rbg[i] - s.xp[1 + i]; s.F[8 + i] = b[i] - s.xp[4 + i]; }
    for (int i = 0; i < 3; ++i) {
      s.dFdx1[i * 6 + i] = ca;
      s.dFdx1[3 * 6 + i] = (float)(2.0 * (double)g[i]);
      for (int j = 0; j < 3; ++j) s.dFdx1[(5 + i) * 6 + j] = Rb[i * 3 + j];
    }
    s.dFdx1[5 * 6 + 4] = Rbg[2];  s.dFdx1[5 * 6 + 5] = -Rbg[1];
    s.dFdx1[6 * 6 + 3] = -Rbg[2]; s.dFdx1[6 * 6 + 5] = Rbg[0];
    s.dFdx1[7 * 6 + 3] = Rbg[1];  s.dFdx1[7 * 6 + 4] = -Rbg[0];
    const double k2 = 2.0 * (double)sa * (double)ca;
    for (int e = 0; e < 9; ++e) {
      s.Pz[e] = sa * sa * s.Rv[e] + ca * ca * Rs[e];
      s.dP3[e] = (float)(k2 * (double)(s.Rv[e] - Rs[e]));
    }
  });
  chol_inverse(s.W3, s.Pz, s.ch, 3);
  each(9, [&](int e) {
    const int i = e / 3, j = e % 3;
    s.Wm[i * 11 + j] = s.W3[e];
    s.T3[e] = (-s.W3[i * 3] * s.dP3[j] + -s.W3[i * 3 + 1] * s.dP3[3 + j]) + -s.W3[i * 3 + 2] * s.dP3[6 + j];   // (-W)*dPda
  });
  each(9, [&](int e) {
    const int i = e / 3, j = e % 3;
    s.dW3[e] = (s.T3[i * 3] * s.W3[j] + s.T3[i * 3 + 1] * s.W3[3 + j]) + s.T3[i * 3 + 2] * s.W3[6 + j];        // ... *W
  });
  mm<6, 11, 11>(s.M6a, s.dFdx1, 1, 6, s.Wm, 11, 1);           // dFdx1.T()*W   (6x11)
  // lanes 0..5: one element of the column 0.5*dFdx1.T()*dWda*F + dFdx1.T()*W*dFda, and of dFdx1.T()*W*F
  // lane 8: JtJ(0,0) and JtF[0]
  {
    const int l = lane_id();
    if (l < 6) {
      double r[3];
      for (int j = 0; j < 3; ++j) {                            // row l of (0.5*dFdx1.T())*dWda: terms q = 0..2 (the rest are zeros)
        double t = (0.5 * (double)s.dFdx1[l]) * (double)s.dW3[j];
        t += (0.5 * (double)s.dFdx1[6 + l]) * (double)s.dW3[3 + j];
        t += (0.5 * (double)s.dFdx1[12 + l]) * (double)s.dW3[6 + j];
        r[j] = t;
      }
      double cd = r[0] * (double)s.F[0];
      cd += r[1] * (double)s.F[1];
      cd += r[2] * (double)s.F[2];
      float cf = s.M6a[l * 11] * s.dFda[0], jf = s.M6a[l * 11] * s.F[0];
      for (int q = 1; q < 11; ++q) { cf += s.M6a[l * 11 + q] * s.dFda[q]; jf += s.M6a[l * 11 + q] * s.F[q]; }
      const float c = (float)(cd + (double)cf);
      s.JtJ[(1 + l) * 7] = c;
      s.JtJ[1 + l] = c;
      s.JtF[1 + l] = jf;
    } else if (l == 8) {
      double v0[3], v1[3];
      // 0.25*F*dWda*P*dWda*F
      for (int j = 0; j < 3; ++j) {
        double t = (0.25 * (double)s.F[0]) * (double)s.dW3[j];
        t += (0.25 * (double)s.F[1]) * (double)s.dW3[3 + j];
        t += (0.25 * (double)s.F[2]) * (double)s.dW3[6 + j];
        v0[j] = t;
      }
      for (int j = 0; j < 3; ++j) { double t = v0[0] * (double)s.Pz[j]; t += v0[1] * (double)s.Pz[3 + j]; t += v0[2] * (double)s.Pz[6 + j]; v1[j] = t; }
      for (int j = 0; j < 3; ++j) { double t = v1[0] * (double)s.dW3[j]; t += v1[1] * (double)s.dW3[3 + j]; t += v1[2] * (double)s.dW3[6 + j]; v0[j] = t; }
      double t0 = v0[0] * (double)s.F[0];
      t0 += v0[1] * (double)s.F[1];
      t0 += v0[2] * (double)s.F[2];
      // dFda*dWda*F
      float u[3];
      for (int j = 0; j < 3; ++j) u[j] = (s.dFda[0] * s.dW3[j] + s.dFda[1] * s.dW3[3 + j]) + s.dFda[2] * s.dW3[6 + j];
      const float t1 = (u[0] * s.F[0] + u[1] * s.F[1]) + u[2] * s.F[2];
      // dFda*W: entries 0..2 from W3, entry 3 zero, entries 4..10 = row 0 of W7 (dFda[4] = 1)
      float m[11];
      for (int j = 0; j < 3; ++j) m[j] = (s.dFda[0] * s.W3[j] + s.dFda[1] * s.W3[3 + j]) + s.dFda[2] * s.W3[6 + j];
      m[3] = 0.0f;
      for (int j = 0; j < 7; ++j) m[4 + j] = s.W7[j];
      const float t2 = ((m[0] * s.dFda[0] + m[1] * s.dFda[1]) + m[2] * s.dFda[2]) + m[4];
      float t3 = m[0] * s.F[0];
      for (int j = 1; j < 11; ++j) t3 += m[j] * s.F[j];
      s.JtJ[0] = (float)((t0 + (double)t1) + (double)t2);
      // 0.5*F*dWda*F + dFda*W*F
      for (int j = 0; j < 3; ++j) {
        double t = (0.5 * (double)s.F[0]) * (double)s.dW3[j];
        t += (0.5 * (double)s.F[1]) * (double)s.dW3[3 + j];
        t += (0.5 * (double)s.F[2]) * (double)s.dW3[6 + j];
        v1[j] = t;
      }
      double t4 = v1[0] * (double)s.F[0];
      t4 += v1[1] * (double)s.F[1];
      t4 += v1[2] * (double)s.F[2];
      s.JtF[0] = (float)(t4 + (double)t3);
    }
    __syncwarp();
  }
  mm<6, 11, 6>(s.M6b, s.M6a, 11, 1, s.dFdx1, 6, 1);           // (dFdx1.T()*W)*dFdx1
  each(36, [&](int e) { s.JtJ[(1 + e / 6) * 7 + 1 + e % 6] = s.M6b[e]; });
}

__device__ __forceinline__ float saturate_f(float t, float limit) { return (t > limit) ? limit : ((t < -limit) ? -limit : t); }

// ---- rebvio.cpp:183-233 and 255-273; one warp per lane ----
// vg / wx / x: the tracker's results for lane b at vg + b*vg_stride etc. (LMState::vel and lane_out in the fused call, a
// staging buffer in the stage call). Writes the second half's inputs into lane_in and the per-frame record into out.
__global__ void __launch_bounds__(32) k_imu_fuse(ImuLane* __restrict__ st, const rbv_imu_in* __restrict__ in, rbv_imu_cfg cfg,
                                                 const float* __restrict__ vg, int vg_stride, const float* __restrict__ wx,
                                                 int wx_stride, const float* __restrict__ x, int x_stride,
                                                 float* __restrict__ lane_in, float max_radius, rbv_imu_out* __restrict__ out) {
  __shared__ FusionSh s;
  __shared__ float Rs3[9], sRGyro[9], sRGBias[9];
  const int b = blockIdx.x;
  ImuLane& S = st[b];
  const float frame_dt = in[b].frame_dt;
  each(36, [&](int e) { s.Wx[e] = wx[(size_t)b * wx_stride + e]; });
  each(6, [&](int e) { s.X[e] = x[(size_t)b * x_stride + e]; });
  each(3, [&](int e) { s.Vg[e] = vg[(size_t)b * vg_stride + e]; });
  each(9, [&](int e) { s.R[e] = S.R[e]; s.Wb[e] = S.W_Bg[e]; });

  // ---- Core::gyroBiasCorrection (core.cpp:264-287) with the noise matrices of rebvio.cpp:185-186
  one([&]() {
    const float gb = cfg.gyro_bias_std_dev * cfg.gyro_bias_std_dev * frame_dt * frame_dt;
    const float gg = cfg.gyro_std_dev * cfg.gyro_std_dev * frame_dt * frame_dt;
    for (int i = 0; i < 9; ++i) { sRGBias[i] = (i % 4 == 0) ? gb : 0.0f; sRGyro[i] = (i % 4 == 0) ? gg : 0.0f; }
    float t[9], u[9];
    invert3(s.Wg, sRGyro);
    invert3(t, s.Wb);
    for (int i = 0; i < 9; ++i) t[i] += sRGBias[i];
    invert3(s.Wb, t);                                    // _Wb = invert(invert(_Wb) + _Rb)
    for (int i = 0; i < 9; ++i) t[i] = s.Wg[i] + s.Wb[i];
    invert3(s.iWgWb, t);
    mat3_mul(t, s.iWgWb, s.Wg);
    for (int i = 0; i < 9; ++i) t[i] = ((i % 4 == 0) ? 1.0f : 0.0f) - t[i];
    mat3_mul(u, s.Wg, t);                                // Wg*(I - iWgWb*Wg)
    for (int i = 0; i < 36; ++i) s.Wxb[i] = s.Wx[i];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) s.Wxb[(3 + i) * 6 + 3 + j] += u[i * 3 + j];
  });
  mm(s.X1, s.Wx, 6, 1, s.X, 1, 0, 6, 6, 1);              // X1 = _Wx*_X  (+ a zero vector, core.cpp:279)
  chol_inverse(s.inv6, s.Wxb, s.ch, 6);
  mm(s.X, s.inv6, 6, 1, s.X1, 1, 0, 6, 6, 1);            // _X = Cholesky(Wxb)^-1 * X1
  one([&]() {
    float t[3], d[3];
    mat3_vec(t, s.Wg, s.X + 3);
    mat3_vec(d, s.iWgWb, t);                             // dgbias = iWgWb*(Wg*_X[3:6] + _Wb*0)
    for (int i = 0; i < 3; ++i) S.Bg[i] += d[i];
    for (int i = 0; i < 9; ++i) { S.W_Bg[i] = s.Wg[i] + s.Wb[i]; S.RGBias[i] = sRGBias[i]; }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) s.Wx[(3 + i) * 6 + 3 + j] += s.Wg[i * 3 + j];
    // ---- rebvio.cpp:191-197: Rgva = R; R.T() = R0*R.T(); Vgv = R0*Vg + dVgv
    for (int i = 0; i < 9; ++i) s.Rgva[i] = s.R[i];
    so3_exp(s.R0, s.X + 3);
    float rt[9];
    mat3_mul_bt(rt, s.R0, s.R);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) s.R[i * 3 + j] = rt[j * 3 + i];
    mat3_vec(t, s.R0, s.Vg);
    for (int i = 0; i < 3; ++i) s.V[i] = t[i] + s.X[i];
  });
  chol_inverse(s.inv6, s.Wx, s.ch, 6);                   // R_Xgv
  one([&]() {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) { s.P_V[i * 3 + j] = s.inv6[i * 6 + j]; s.P_W[i * 3 + j] = s.inv6[(3 + i) * 6 + 3 + j]; }
    // ---- Core::estimateLs4Acceleration(-Vgv/frame_dt, Av, R, frame_dt) (core.cpp:289-336)
    float vel[3], t[3];
    for (int i = 0; i < 3; ++i) vel[i] = -s.V[i] / frame_dt;
    for (int k = 4; k >= 1; --k) { mat3t_vec(t, s.R, S.lsV[k - 1]); for (int i = 0; i < 3; ++i) S.lsV[k][i] = t[i]; }
    for (int i = 0; i < 3; ++i) S.lsV[0][i] = vel[i];
    for (int i = 0; i < 3; ++i) S.lsDt[i] = S.lsDt[i + 1];
    S.lsDt[3] = frame_dt;
    float T[5];
    T[0] = 0.0f;
    float mt = 0.0f;
    for (int i = 0; i < 4; ++i) { T[i + 1] = T[i] + S.lsDt[i]; mt += T[i + 1]; }
    mt = (float)((double)mt / 5.0);
    float den = 0.0f;
    for (int i = 0; i < 5; ++i) den += (T[i] - mt) * (T[i] - mt);
    // `V[3]` in core.cpp:323 reads one float past the static Vector3f V; in the reference as built here that is zero
    // padding of .bss (checked with nm on oracle/_ref), so the fifth term of the mean is 0, not V3[i].
    const float v_oob = 0.0f;
    for (int i = 0; i < 3; ++i) {
      const float vm = (float)((double)(S.lsV[0][i] + S.lsV[1][i] + S.lsV[2][i] + S.lsV[3][i] + v_oob) / 5.0);
      float num = (S.lsV[0][i] - vm) * (T[4] - mt);
      num += (S.lsV[1][i] - vm) * (T[3] - mt);
      num += (S.lsV[2][i] - vm) * (T[2] - mt);
      num += (S.lsV[3][i] - vm) * (T[1] - mt);
      num += (S.lsV[4][i] - vm) * (T[0] - mt);
      if (den > 0.0) S.Av[i] = num / den;
    }
    // ---- Core::estimateMeanAcceleration(cacc, As, R) (core.cpp:338-352)
    for (int k = 3; k >= 1; --k) { mat3t_vec(t, s.R, S.maA[k - 1]); for (int i = 0; i < 3; ++i) S.maA[k][i] = t[i]; }
    for (int i = 0; i < 3; ++i) S.maA[0][i] = in[b].cacc[i];
    for (int i = 0; i < 3; ++i) S.As[i] = (float)(0.25 * (double)(S.maA[0][i] + S.maA[1][i] + S.maA[2][i] + S.maA[3][i]));
    // rebvio.cpp:206-209
    const float dt4 = frame_dt * frame_dt * frame_dt * frame_dt;
    for (int i = 0; i < 9; ++i) s.Rv[i] = s.P_V[i] / dt4;
    for (int i = 0; i < 6; ++i) s.Xgva[i] = s.X[i];
    for (int i = 0; i < 3; ++i) { s.as[i] = S.As[i]; s.av[i] = S.Av[i]; }
    s.K = S.K;
    s.stopped = 0;
  });

  const bool with_acc = S.num_frames > 4 + cfg.init_bias_frame_num;   // warp-uniform
  if (with_acc) {
    // ---- Core::estimateBias (core.cpp:355-414); constants of SABEstimator::State (sab_estimator.hpp:56-71)
    each(7, [&](int e) { s.Xs[e] = S.X[e]; });
    each(49, [&](int e) { s.T7[e] = S.P[e]; s.F7[e] = 0.0f; s.Q[e] = 0.0f; });
    one([&]() {
      s.G = cfg.g_norm;
      s.Rg = cfg.g_norm_uncertainty * cfg.g_norm_uncertainty;
      const float rs = cfg.acc_std_dev * cfg.acc_std_dev, qg = cfg.g_uncertainty * cfg.g_uncertainty;
      const float qb = cfg.vbias_std_dev * cfg.vbias_std_dev;
      for (int i = 0; i < 9; ++i) Rs3[i] = (i % 4 == 0) ? rs : 0.0f;
      s.F7[0] = 1.0f;
      for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) s.F7[(1 + i) * 7 + 1 + j] = s.R[j * 3 + i];   // Rot.T()
        s.F7[(4 + i) * 7 + 4 + i] = 1.0f;
      }
      const float* g = s.Xs + 1;
      const float GP[9] = {0.0f, g[2], -g[1], -g[2], 0.0f, g[0], g[1], -g[0], 0.0f};
      const float tn = tanf(s.Xs[0]);
      s.Q[0] = (float)((double)S.P_Kp / (1.0 + (double)(tn * tn)));
      float t[9], u[9];
      for (int i = 0; i < 3; ++i)                                                  // GProd.T()*Qrot
        for (int j = 0; j < 3; ++j) t[i * 3 + j] = GP[i] * s.P_W[j] + GP[3 + i] * s.P_W[3 + j] + GP[6 + i] * s.P_W[6 + j];
      mat3_mul(u, t, GP);
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
          s.Q[(1 + i) * 7 + 1 + j] = u[i * 3 + j] + ((i == j) ? qg : 0.0f);
          s.Q[(4 + i) * 7 + 4 + j] = (i == j) ? qb : 0.0f;
        }
    });
    mm(s.xp, s.F7, 7, 1, s.Xs, 1, 0, 7, 7, 1);             // _X = F*_X
    mm(s.T7b, s.F7, 7, 1, s.T7, 7, 1, 7, 7, 7);            // F*_P
    mm(s.Pp, s.T7b, 7, 1, s.F7, 1, 7, 7, 7, 7);            // (F*_P)*F.T()
    each(49, [&](int e) { s.Pp[e] += s.Q[e]; });
    each(7, [&](int e) { s.Xs[e] = s.xp[e]; });
    chol_inverse(s.W7, s.Pp, s.ch, 7);
    each(121, [&](int e) {
      const int i = e / 11, j = e % 11;
      float w = 0.0f;
      if (i == 3 && j == 3) w = (float)(1.0 / (double)s.Rg);
      if (i >= 4 && j >= 4) w = s.W7[(i - 4) * 7 + j - 4];
      s.Wm[e] = w;
    });
    each(66, [&](int e) { const int i = e / 6, j = e % 6; s.dFdx1[e] = (i >= 8 && j == i - 5) ? 1.0f : 0.0f; });
    // ---- SABEstimator::gaussNewton(_X, 20) with a_tol = r_tol = 0: never leaves early (sab_estimator.cpp:21-42). An
    // iteration that leaves X bit-identical makes every later one a repeat of itself, so stopping there changes nothing.
    double jv[8];                    // row `lane` of the eigenvector matrix, carried from one iteration to the next
    bool warm = false, try_ldl = true;   // once the conditioning test has failed the later iterations (nearly the same matrix) go straight to the eigen-solve
    for (int it = 0; it < 20; ++it) {
      sab_problem(s, Rs3);
      each(49, [&](int e) { const int i = e / 7, j = e % 7; s.S7[e] = 0.5 * ((double)s.JtJ[e] + (double)s.JtJ[j * 7 + i]); });
      each(7, [&](int e) { s.rhs[e] = -(double)s.JtF[e]; });
      double h;
      if (!(try_ldl && solve_ldl_warp<7>(s.S7, s.rhs, h))) {
        try_ldl = false;
        each(64, [&](int e) { const int i = e >> 3, j = e & 7; s.A8[e] = (i < 7 && j < 7) ? s.S7[i * 7 + j] : 0.0; });
        h = solve_jacobi8_warp(s.A8, s.rhs, jv, warm, s.V8, s.T8);
        warm = true;
      }
      const int l = lane_id();
      float xn = 0.0f;
      if (l < 7) {
        xn = (float)((double)s.Xs[l] + h);
        if (l == 0) xn = atan2f(sinf(xn), cosf(xn));
        if (l >= 4) xn = saturate_f(xn, (float)(5e-1 / 25));
      }
      const unsigned changed = __ballot_sync(FULL, l < 7 && __float_as_uint(xn) != __float_as_uint(s.Xs[l]));
      __syncwarp();
      if (l < 7) s.Xs[l] = xn;
      __syncwarp();
      if (!changed) break;
    }
    sab_problem(s, Rs3);
    chol_inverse(s.T7, s.JtJ, s.ch, 7);                    // _P
    each(49, [&](int e) { S.P[e] = s.T7[e]; });
    each(7, [&](int e) { S.X[e] = s.Xs[e]; });
    one([&]() {
      float k = tanf(s.Xs[0]);
      if (k < 0 || isnan(k) || isinf(k)) k = 0;
      s.K = k;
      for (int i = 0; i < 3; ++i) { S.g_est[i] = s.Xs[1 + i]; S.b_est[i] = s.Xs[4 + i]; }
      float WV[9], wc[3], wxc[3];
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) WV[i * 3 + j] = s.JtJ[(4 + i) * 7 + 4 + j];
      for (int i = 0; i < 3; ++i) wc[i] = s.Xgva[3 + i] - s.Xs[4 + i];
      mat3_vec(wxc, WV, wc);
      for (int i = 0; i < 36; ++i) s.Wxb[i] = s.Wx[i];     // Wb + _Wvw (Wb is zero outside the rotation block: 0 + w)
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) s.Wxb[(3 + i) * 6 + 3 + j] = WV[i * 3 + j] + s.Wx[(3 + i) * 6 + 3 + j];
      for (int i = 0; i < 6; ++i) {                        // _Wvw*_Xvw + WXc
        float a = s.Wx[i * 6] * s.Xgva[0];
        for (int q = 1; q < 6; ++q) a += s.Wx[i * 6 + q] * s.Xgva[q];
        s.X1[i] = a + (i >= 3 ? wxc[i - 3] : 0.0f);
      }
    });
    chol_inverse(s.inv6, s.Wxb, s.ch, 6);
    mm(s.Xgva, s.inv6, 6, 1, s.X1, 1, 0, 6, 6, 1);
  }

  // ---- rebvio.cpp:214-233 composition, :236-241 NaN check, :255-273 pose
  one([&]() {
    float R0u[9];
    if (with_acc) {
      float R0g[9], rt[9], t[3];
      so3_exp(R0g, s.Xgva + 3);
      mat3_mul_bt(rt, R0g, s.Rgva);                        // Rgva.T() = R0gva*Rgva.T()
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) s.Rgva[i * 3 + j] = rt[j * 3 + i];
      mat3_vec(t, R0g, s.Vg);
      for (int i = 0; i < 3; ++i) s.V[i] = t[i] + s.Xgva[i];
      for (int i = 0; i < 9; ++i) R0u[i] = R0g[i];
    } else {
      for (int i = 0; i < 9; ++i) { s.Rgva[i] = s.R[i]; R0u[i] = s.R0[i]; }
    }
    float Vgva[3] = {s.V[0], s.V[1], s.V[2]};
    if (isnan(s.V[0]) || isnan(s.V[1]) || isnan(s.V[2])) {
      for (int i = 0; i < 9; ++i) s.P_V[i] = (i % 4 == 0) ? 3.402823466e+38f : 0.0f;
      for (int i = 0; i < 3; ++i) s.V[i] = 0.0f;
      S.P_Kp = 3.402823466e+38f;
      s.stopped = 1;
      S.stopped = 1;
    }
    S.K = s.K;
    if (with_acc) {
      float u[3], t[3];
      mat3t_vec(u, s.Rgva, S.u_est);
      const float f = dot3(u, S.g_est) / dot3(S.g_est, S.g_est);
      for (int i = 0; i < 3; ++i) u[i] = u[i] - f * S.g_est[i];
      unit3(u, u);
      for (int i = 0; i < 3; ++i) S.u_est[i] = u[i];
      const float ey[3] = {0.0f, 1.0f, 0.0f}, ex[3] = {1.0f, 0.0f, 0.0f};
      float R1[9], R2[9], Rg[9];
      so3_from_to(R1, S.g_est, ey);
      mat3_vec(t, R1, u);
      so3_from_to(R2, t, ex);
      mat3_mul(Rg, R2, R1);
      for (int i = 0; i < 9; ++i) S.R_global[i] = Rg[i];
      for (int i = 0; i < 3; ++i)                          // Pos += -R_global*Vgva*K
        S.Pos[i] += (-Rg[i * 3] * Vgva[0] + -Rg[i * 3 + 1] * Vgva[1] + -Rg[i * 3 + 2] * Vgva[2]) * s.K;
    }
    rbv_imu_out& O = out[b];
    so3_ln(O.orientation, S.R_global);
    for (int i = 0; i < 3; ++i) { O.position[i] = S.Pos[i]; O.V[i] = s.V[i]; O.Bg[i] = S.Bg[i]; O.Av[i] = S.Av[i]; O.As[i] = S.As[i]; }
    for (int i = 0; i < 9; ++i) { O.R[i] = s.R[i]; O.R0[i] = R0u[i]; O.Rgva[i] = s.Rgva[i]; O.P_V[i] = s.P_V[i]; }
    for (int i = 0; i < 6; ++i) { O.Xgv[i] = s.X[i]; O.Xgva[i] = s.Xgva[i]; }
    for (int i = 0; i < 7; ++i) O.X[i] = S.X[i];
    O.K = s.K;
    O.initialized = S.initialized;
    O.stopped = s.stopped;
    O.num_frames = ++S.num_frames;
    float* li = lane_in + (size_t)b * IN_STRIDE;           // the second half reads its arguments here
    for (int i = 0; i < 9; ++i) { li[IN_R + i] = R0u[i]; li[IN_RVEL + i] = s.P_V[i]; li[IN_RBACK + i] = s.Rgva[i]; }
    for (int i = 0; i < 3; ++i) li[IN_VEL + i] = s.V[i];
    li[IN_MAXRADIUS] = max_radius;
  });
}

}  // namespace

// ---- launchers (entry points in rbv_api.cu) ----
int fu_reset(rbv_ctx* c, const rbv_imu_cfg* cfg) {
  const int B = c->P.lanes;
  if (!c->imu_state) {
    RBV_CUDA(cudaMalloc(&c->imu_state, sizeof(ImuLane) * (size_t)B));
    RBV_CUDA(cudaMalloc(&c->imu_in, sizeof(rbv_imu_in) * (size_t)B));
    RBV_CUDA(cudaMalloc(&c->imu_out, sizeof(rbv_imu_out) * (size_t)B));
    RBV_CUDA(cudaMalloc(&c->imu_stage, sizeof(float) * 48 * (size_t)B));
    RBV_CUDA(cudaMallocHost(&c->imu_host, (sizeof(rbv_imu_in) + sizeof(rbv_imu_out) + sizeof(float) * 48) * (size_t)B));
  }
  c->imu_cfg = *cfg;
  // a freshly constructed Rebvio: types::ImuState (imu.hpp:170-187), SABEstimator::State (sab_estimator.hpp:56-71),
  // locals of stateEstimationProcess (rebvio.cpp:95-115)
  ImuLane L;
  memset(&L, 0, sizeof(L));
  for (int i = 0; i < 9; i += 4) { L.RGBias[i] = 1.0f; L.W_Bg[i] = 1.0f / 100.0f; L.R[i] = 1.0f; L.R_global[i] = 1.0f; }
  L.u_est[0] = 1.0f;
  L.X[0] = (float)M_PI_4;
  L.X[2] = cfg->g_norm;
  L.P[0] = cfg->scale_stdd_dev_init * cfg->scale_stdd_dev_init;
  for (int i = 1; i < 4; ++i) L.P[i * 8] = 100.0f;
  for (int i = 4; i < 7; ++i) L.P[i * 8] = (float)((double)(cfg->vbias_std_dev * cfg->vbias_std_dev) * 1e1);
  L.K = 1.0f;
  L.P_Kp = 5e-6f;
  ImuLane* h = new ImuLane[B];
  for (int b = 0; b < B; ++b) h[b] = L;
  cudaError_t e = cudaMemcpyAsync(c->imu_state, h, sizeof(ImuLane) * (size_t)B, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  delete[] h;
  RBV_CUDA(e);
  return RBV_OK;
}

int fu_pre(rbv_ctx* c) {
  const int B = c->P.lanes;
  RBV_LAUNCH(c, k_imu_pre, ((B + 31) / 32), (32), 0, static_cast<ImuLane*>(c->imu_state),
             static_cast<const rbv_imu_in*>(c->imu_in), c->imu_cfg, c->lane_in, static_cast<rbv_imu_out*>(c->imu_out), B);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

int fu_fuse(rbv_ctx* c, const float* vg, int vg_stride, const float* wx, int wx_stride, const float* x, int x_stride,
            float max_radius) {
  RBV_LAUNCH(c, k_imu_fuse, (c->P.lanes), (32), 0, static_cast<ImuLane*>(c->imu_state),
             static_cast<const rbv_imu_in*>(c->imu_in), c->imu_cfg, vg, vg_stride, wx, wx_stride, x, x_stride, c->lane_in,
             max_radius, static_cast<rbv_imu_out*>(c->imu_out));
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}
