// rbv_internal.cuh — shared declarations of the B200-native REBVIO hot path (sm_100a).
//
// Layout in HBM (all arrays lane-major, B = lanes = independent image sequences processed per launch):
//   images      float  [B][rows][cols]        input frame, scale-space ping/pong buffers, row-prefix scratch
//   id mask     int32  [B][rows*cols]         per map: pixel -> keyline id (-1 none)  (EdgeMap::mask(), edge_map.hpp:130)
//   field       uint32 [B][rows*cols]         packed distance field (|r| << 24 | 0xFFFFFF - id), 0xFFFFFFFF = empty
//   keylines    SoA    [field][B][kmax]       float2 for the 2-vectors, float / int32 / uint32 scalars
// Everything is float32 evaluated in the reference's order; the translation units are compiled with -fmad=false so no
// multiply-add is ever contracted (the reference parity build is -ffp-contract=off).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rebvio_b200.h"

#define RBV_RHO_MAX 20.0f   // types/keyline.hpp:19
#define RBV_RHO_MIN 1e-3f   // types/keyline.hpp:20
#define RBV_RHO_INIT 1.0f   // types/keyline.hpp:21
#define RBV_FIELD_EMPTY 0xFFFFFFFFu
#define RBV_TUNE_BINS 100
#define RBV_MAX_QBINS 1024
#define RBV_MAX_LM_ITERS 15

// ---- keyline SoA of one B-lane map (device pointers, each [B][kmax]) ----
struct KeylineSoA {
  float2* pos;            // KeyLine::pos
  float2* pos_img;        // KeyLine::pos_img
  float2* match_pos_img;  // KeyLine::match_pos_img
  float2* grad;           // KeyLine::gradient
  float2* match_grad;     // KeyLine::match_gradient
  float* gnorm;           // gradient_norm
  float* match_gnorm;     // match_gradient_norm
  float* rho;
  float* srho;            // sigma_rho
  int* id_prev;
  int* id_next;
  int* match_id;
  int* match_fwd;         // match_id_forward
  int* match_kf;          // match_id_keyframe
  unsigned* matches;
  unsigned* pixel;        // detection pixel (key of EdgeMap::mask())
  float* residual;        // Core::minimizeVel's residuals[] (core.cpp:158), kept with the map
};

struct rbv_map {
  KeylineSoA k;
  int* mask;         // [B][npix]
  int* count;        // [B]
  float* threshold;  // [B]  EdgeMap::threshold_
  uint64_t* ts_us;   // host copy only
  void* slab;        // one allocation backing everything above
  size_t slab_bytes;
  cudaEvent_t ready; // recorded behind a detect on the front-end stream; tracker entries wait for it (fe_pending)
  int fe_pending, fe_valid;
  unsigned* field;   // [B][N] this map's own distance field and [B][kmax][2] packed targets, built behind its detect on the front-end
  float4* tgt;       // stream (RBV_TUNE_FIELD_IN_FE); nullptr until first needed. field_built: valid for the keylines now in the map
  int field_built;
  cudaEvent_t used;  // recorded on the tracker stream behind every tracker / mirror entry that was given this map; a detect INTO the map waits for it
  int used_valid;
  int* h_fe;         // pinned [B] keyline counts + [B] auto thresholds of that detect (rbv_detect_wait)
};

// Levenberg-Marquardt state of Core::minimizeVel, one per lane, device resident
struct LMState {
  float vel[3];
  float F;
  float JtJ[9];
  float JtF[3];
  float u, v;
  float vnew[3];
  float h[3];
  float sigma_rho_min;
  float Rvel[9];
  float trace[(RBV_MAX_LM_ITERS + 1) * 16];
};

struct DevParams {  // constant per context
  float fm, cx, cy;
  int rows, cols, npix, kmax, lanes;
  // detector
  float pn_threshold;
  float dog_threshold;
  int keylines_max;
  // map cfg
  float pixel_uncertainty_match, match_threshold_norm, cang_min_edge, regularization_threshold;
  // core cfg
  float search_range, reweight_distance, match_treshold, pixel_uncertainty, quantile_cutoff, reshape_q_abs;
  int quantile_num_bins, iterations, field_r0, field_nr;
  unsigned min_matches;
  float pinv[75];
};

// staging slots for uploaded frames: the one the front end reads, the one being uploaded, and one more because the front end may lag
// a frame behind the host when it runs ahead of the tracker (RBV_TUNE_FE_STREAM)
#define RBV_RAW_SLOTS 3
struct rbv_ctx {
  DevParams P;
  rbv_camera cam;
  rbv_detector_cfg det;
  rbv_map_cfg mapc;
  rbv_core_cfg core;
  int device;
  int reduce_mode;
  cudaStream_t stream;
  bool own_stream;
  int widths[2][3];
  // images
  float* img;       // [B][N]
  float* rowsum;    // [2][B][N] row-prefix scratch per scale
  float* bufA;      // [2][B][N]
  float* bufB;      // [2][B][N]
  // detect scratch
  unsigned* flags;  // [B][rows][wpr] bit flags
  int* rowcount;    // [B][rows]
  int* rowoffset;   // [B][rows]
  float* thr_dev;   // [B] servoed thresholds of this detect = config_->threshold of every lane (persistent: k_servo advances it)
  int* last_count;  // [B] keyline counts of the most recent detect (EdgeDetector::keylines_count_), input of k_servo
  // tracker scratch
  unsigned* field;  // [B][N]
  float4* tgt;      // [B][kmax][2] packed target keylines of the distance-field map (k_pack_targets)
  rbv_map* field_map;
  float* prod;      // [27][B][kmax] (+ a chunk of slack) ORDERED mode: the 10 float32 addends of tryVel / the 27 of extRotVel per keyline
  float* fiabs;     // [B][kmax]
  int* fwd_head;    // [B][kmax]
  int* fwd_next;    // [B][kmax]
  float* reg_r;     // [B][kmax]
  float* reg_s;     // [B][kmax]
  LMState* lm;      // [B]
  float* lane_in;   // [B][64] per-call lane inputs (R, vel, Rvel, Rback ...)
  float* lane_out;  // [B][64] per-call lane outputs
  int* lane_cnt;    // [B][8] counters (matches, kf, regularised ...)
  unsigned long long* probes;  // [B]
  // FAST-mode reductions: per-block FP64 partials, per-lane tickets, per-block "last matched |fi|" (ping-pong)
  double* partials;  // [B][nblk][27]
  unsigned* ticket;  // [B]
  int* lm_done;      // [B] k_minimize_ord*: finished LM calls of the lane in the running launch (reset by k_quantile)
  int* ord_flags;    // [B][nblk] ORDERED: evaluation block j of the running k_tryvel_ord* launch has left its addends (set by the block, cleared by the chain's loader)
  float* blkagg;     // [2][B][nblk]
  int nblk;          // blocks of 256 keylines per lane
  int tv_parity;
  // pinned host staging
  float* h_in;
  float* h_out;
  int* h_cnt;
  float* h_lm;
  int* h_misc;        // [2][B] counts / thresholds of the small synchronous reads (rbv_map_size, rbv_map_threshold, detect outputs)
  void* h_trk;        // rbv_track_imu's own result block (see trk_block_bytes, rbv_api.cu)
  cudaEvent_t ev_upload;   // behind the H2D of a detect call's host frames
  // frame upload: double-buffered raw staging on a copy stream (rbv_prefetch_frames) so that the H2D of frame k+1 overlaps
  // the tracker of frame k; u8 frames are converted (x 3.0, rebvio.cpp:43) on the device
  cudaStream_t copy_stream;
  cudaStream_t fe_stream;     // front-end stream (RBV_TUNE_FE_STREAM): rbv_detect* next to the tracker of the previous frame pair
  cudaEvent_t ev_tracker;
  cudaStream_t side_stream;   // second branch of rbv_track_g1: rotate + quantile of the old map next to the field build of the new one
  cudaEvent_t ev_fork, ev_join;
  void* raw[RBV_RAW_SLOTS];               // [B][N] bytes (u8) or floats (f32) as uploaded
  cudaEvent_t raw_ready[RBV_RAW_SLOTS];   // upload of slot i finished (recorded on copy_stream)
  cudaEvent_t raw_free[RBV_RAW_SLOTS];    // the front end has consumed slot i (recorded on stream)
  const void* raw_key[RBV_RAW_SLOTS];     // images[0] of the prefetch that filled slot i (nullptr = empty)
  int raw_u8[RBV_RAW_SLOTS];
  float* rawf[RBV_RAW_SLOTS];  // [B][N] float32 x 3 of a prefetched u8 slot, converted on the copy stream (off the front-end chain); lazily allocated
  int raw_conv[RBV_RAW_SLOTS]; // rawf[i] holds the converted frames of slot i
  int raw_next;
  uint2* undist_map;          // [N] fixed-point remap of cv::undistort (rbv_set_undistort), nullptr = frames arrive undistorted
  unsigned long long launches;  // kernels launched since creation
  unsigned long long klaunch[64];   // per kernel id
  int tune[RBV_TUNE_COUNT];         // rbv_set_tuning: implementation choices (0 = automatic)
  int prof_mode;                    // -1 off, -2 every kernel, >= 0 that kernel id only
  void* prof;                       // ProfState* (rbv_api.cu)
  int seq_fused;                    // inside rbv_track_g1/g2: k_rotate zeroes the lane counters, the stages skip k_zero_cnt
  rbv_map* seq_new;                 // rbv_track_g2: map whose keyline count / auto threshold ride home in lane_cnt slots 4, 5
  const float* seq_resolve_agg;     // rbv_track_g1: block aggregates of the last tryVel call, for k_fwd_link to clear the residual markers
  int seq_resolve_kpb;
  int seq_reg_pending;              // rbv_track_g2: regularize1Iter's write phase is pending and will be done by k_kalman
  int seq_min_matches;              // rbv_track_g2: lanes whose directed-match count is below this skip regularise + Kalman (0 = off)
  void* mtx;                        // std::recursive_mutex*: every C-ABI entry locks its context (the reference calls detect and the
                                    // tracker stages from two threads, rebvio.cpp:28-29)
  void *imu_state, *imu_in, *imu_out, *imu_host;   // IMU fusion block (rbv_fusion.cu): per-lane ImuLane, per-frame rbv_imu_in / rbv_imu_out, pinned mirror
  float* imu_stage;                 // [B][48] tracker results handed to the stage call rbv_imu_fuse (Vg 3, Wx 36, X 6)
  rbv_imu_cfg imu_cfg;
  void* graphs;                     // GraphCache* (rbv_api.cu): captured stage sequences, see run_graphed
};

// lane_in slots
enum { IN_R = 0, IN_VEL = 9, IN_RVEL = 12, IN_RBACK = 21, IN_SRMIN = 30, IN_MAXRADIUS = 31, IN_STRIDE = 64 };
// lane_out slots
enum { OUT_JTJ = 0, OUT_JTF = 9, OUT_SCORE = 12, OUT_WX = 16, OUT_X = 52, OUT_OK = 58, OUT_Q = 59, OUT_STRIDE = 64 };
// lane_cnt slots
enum { CNT_FWD = 0, CNT_DIR = 1, CNT_KF = 2, CNT_REG = 3, CNT_NKEY = 4, CNT_ATHR = 5, CNT_STRIDE = 8 };

// ---- kernel registry: every launch goes through RBV_LAUNCH so that launches are counted per kernel and can be
// bracketed with CUDA events on the launching stream (rbv_profile_*; bench.py's roofline numbers come from there) ----
#define RBV_KERNEL_LIST(X)                                                                                         \
  X(k_fill) X(k_rowscan) X(k_colbox) X(k_detect_flags) X(k_detect_scan) X(k_detect_emit) X(k_join) X(k_tune)       \
  X(k_dog_mag) X(k_field_build) X(k_rotate) X(k_quantile) X(k_tryvel_eval) X(k_tryvel_reduce) X(k_zero_cnt)        \
  X(k_fwd_link) X(k_fwd_resolve) X(k_extrot_rows) X(k_extrot_reduce) X(k_directed) X(k_reg_eval) X(k_reg_apply)    \
  X(k_kalman) X(k_lm_init) X(k_tryvel_fast) X(k_residual_resolve) X(k_extrot_fast)        \
  X(k_field_band) X(k_boxpass) X(k_u8_to_f32) X(k_detect_emit_wide) X(k_pack_targets) X(k_rowscan4) \
  X(k_tryvel_fast2) X(k_tryvel_fast4) X(k_rowscan_lat) X(k_colbox_lat) X(k_undistort_map) X(k_undistort) \
  X(k_imu_pre) X(k_imu_fuse) X(k_servo) X(k_tryvel_ord2) X(k_tryvel_ord4) X(k_extrot_ord) X(k_boxpass2) X(k_chain_tryvel) X(k_chain_extrot) X(k_minimize_ord2) X(k_minimize_ord4)
enum {
#define X(n) KID_##n,
  RBV_KERNEL_LIST(X)
#undef X
  KID_COUNT
};
static_assert(KID_COUNT <= 64, "rbv_ctx::klaunch holds 64 kernel ids");
void rbv_prof_begin(rbv_ctx* c, int kid);
void rbv_prof_end(rbv_ctx* c, int kid);
#define RBV_LAUNCH(c, kern, grid, block, smem, ...)                     \
  do {                                                                  \
    rbv_prof_begin((c), KID_##kern);                                    \
    kern<<<dim3 grid, dim3 block, (smem), (c)->stream>>>(__VA_ARGS__);  \
    rbv_prof_end((c), KID_##kern);                                      \
  } while (0)
// device fill of n 32-bit words (replaces cudaMemsetAsync so that every device op is one of our counted kernels)
int rbv_fill32(rbv_ctx* c, void* dst, unsigned value, size_t n_words);

void rbv_set_error(const char* fmt, ...);
#define RBV_CUDA(call)                                                                        \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      rbv_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));    \
      return RBV_ERR_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

// ---- launchers (host functions defined in the .cu files) ----
int fe_init(rbv_ctx* c);   // per-device kernel attributes (dynamic shared memory limits)
int tr_init(rbv_ctx* c);
int fe_scale_space(rbv_ctx* c, const float* d_img, int part);
int fe_uses_fused_box(rbv_ctx* c, const float* d_img);
int fe_detect(rbv_ctx* c, rbv_map* map, float4* tgt = nullptr);   // tgt: also write the packed target records (k_pack_targets' output) of the new keylines
int fe_box_pass_debug(rbv_ctx* c, int scale, int pass, const float* d_in, float* d_out);
int fe_dog_mag_debug(rbv_ctx* c, int lane, int which, float* d_out);

int tr_build_field(rbv_ctx* c, rbv_map* map, bool packed = false);   // packed: the map's target records are already in c->tgt (written by k_detect_emit)
int tr_pack_targets(rbv_ctx* c, rbv_map* map);   // refresh the packed copy after the field map's keylines were replaced
int tr_rotate(rbv_ctx* c, rbv_map* map);                   // R in lane_in[IN_R]
int tr_quantile(rbv_ctx* c, rbv_map* map, float pct, int bins, int to_lm);
int tr_ordered_legacy(const rbv_ctx* c);
int tr_minimize_persist_wanted(const rbv_ctx* c);
int tr_minimize_persist(rbv_ctx* c, rbv_map* map);
size_t tr_ordered_floats(int lanes, int kmax);   // floats of c->prod the chain kernels' addend layout needs
int tr_debug_addends(rbv_ctx* c, int plane, int lane, float* out, int n);
int tr_try_vel(rbv_ctx* c, rbv_map* map, int lm_call);     // lm_call: -1 = single evaluation (vel, srmin in lane_in), t >= 0 = LM call t
int tr_forward_match(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map);
int tr_ext_rot_vel(rbv_ctx* c, rbv_map* new_map, int vel_from_lm);
int tr_directed_match(rbv_ctx* c, rbv_map* new_map, rbv_map* old_map);
int tr_regularize(rbv_ctx* c, rbv_map* map);
int tr_kalman(rbv_ctx* c, rbv_map* map);                   // vel in lane_in[IN_VEL]

// IMU fusion block (rbv_fusion.cu)
int fu_reset(rbv_ctx* c, const rbv_imu_cfg* cfg);
int fu_pre(rbv_ctx* c);                                    // rbv_imu_in in c->imu_in -> lane_in[IN_R], lane_in[IN_VEL] = 0
int fu_fuse(rbv_ctx* c, const float* vg, int vg_stride, const float* wx, int wx_stride, const float* x, int x_stride,
            float max_radius);                             // -> lane_in[IN_R, IN_VEL, IN_RVEL, IN_RBACK, IN_MAXRADIUS], c->imu_out

// ---- device helpers -------------------------------------------------------------------------------------
// 3x3 determinant by Gaussian elimination with partial pivoting (TooN determinant.h:91-146)
__host__ __device__ inline float rbv_det3(const float* in) {
  float A[3][3];
  for (int i = 0; i < 9; ++i) A[i / 3][i % 3] = in[i];
  float det = 1.0f;
  for (int i = 0; i < 3; ++i) {
    int arg = i;
    float best = fabsf(A[i][i]);
    for (int k = i + 1; k < 3; ++k) {
      float v = fabsf(A[k][i]);
      if (v > best) { best = v; arg = k; }
    }
    float pivot = A[arg][i];
    if (arg != i) {
      det *= -1.0f;
      for (int j = i; j < 3; ++j) { float t = A[i][j]; A[i][j] = A[arg][j]; A[arg][j] = t; }
    }
    det *= A[i][i];
    if (det == 0.0f) return 0.0f;
    for (int u = i + 1; u < 3; ++u) {
      float factor = A[u][i] / pivot;
      for (int j = i + 1; j < 3; ++j) A[u][j] = A[u][j] - factor * A[i][j];
    }
  }
  return det;
}

// cofactor inverse / determinant (types/definitions.hpp:40-53)
__host__ __device__ inline void rbv_inv3(const float* m, float* out) {
  float co[9];
  co[0] = m[4] * m[8] - m[5] * m[7];
  co[1] = m[2] * m[7] - m[1] * m[8];
  co[2] = m[1] * m[5] - m[2] * m[4];
  co[3] = m[5] * m[6] - m[3] * m[8];
  co[4] = m[0] * m[8] - m[2] * m[6];
  co[5] = m[2] * m[3] - m[0] * m[5];
  co[6] = m[3] * m[7] - m[4] * m[6];
  co[7] = m[1] * m[6] - m[0] * m[7];
  co[8] = m[0] * m[4] - m[1] * m[3];
  float d = rbv_det3(m);
  for (int i = 0; i < 9; ++i) out[i] = co[i] / d;
}

#ifdef __CUDACC__
// ---- mbarrier + TMA bulk-copy helpers (cp.async.bulk global -> shared, completion on an mbarrier) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// wait on a phase: try_wait with a suspend-time hint parks the warp in hardware instead of spinning (a spinning warp
// takes issue slots from the warps it is waiting for); the loop only re-arms after the hint expires. (Polling with test_wait was
// measured in round 2: no faster - k_boxpass2 230 us either way.)
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  const unsigned a = smem_u32(bar);
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    unsigned ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(a), "r"(parity), "r"(0x989680u) : "memory");
    if (ok) return;
    if ((spin & 255) == 0) {
      if (spin == 0) t0 = clock64();
      else if (clock64() - t0 > 8000000000ll) __trap();   // seconds: a protocol bug must fault, never hang the device
    }
  }
}
// one non-blocking probe of a phase
__device__ __forceinline__ bool mbar_test(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile("{ .reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }


// x86 cvttsd2si / cvttss2si semantics (out of range or NaN -> INT_MIN), which the reference's static_cast<int> compiles to
__device__ __forceinline__ int rbv_d2i(double d) {
  return (d != d || d >= 2147483648.0 || d <= -2147483649.0) ? (-2147483647 - 1) : __double2int_rz(d);
}
__device__ __forceinline__ int rbv_f2i(float f) {
  return (f != f || f >= 2147483648.0f || f < -2147483648.0f) ? (-2147483647 - 1) : __float2int_rz(f);
}
// std::round(float) -> int : half away from zero (getIndex, core.hpp:66-71, edge_map.hpp:119-124)
__device__ __forceinline__ int rbv_pixel_index(int rows, int cols, float row, float col) {
  int r = rbv_f2i(roundf(row)), q = rbv_f2i(roundf(col));
  if (r >= rows || r < 0 || q >= cols || q < 0) return -1;
  return r * cols + q;
}
#endif
