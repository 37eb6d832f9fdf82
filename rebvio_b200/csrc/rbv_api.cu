// rbv_api.cu — C ABI of the B200-native REBVIO hot path (see include/rebvio_b200.h for the reference member function
// each entry point replaces). Host-side setup mirrors FastGaussian::FastGaussian (scale_space.cpp:14-35) and the
// plane-fit pseudo inverse (edge_detector.cpp:55-68). No CPU fallback: without a usable sm_100 device rbv_create fails.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "rbv_internal.cuh"

#define LOCK_CTX(c) std::lock_guard<std::recursive_mutex> ctx_lock_(*static_cast<std::recursive_mutex*>((c)->mtx))
void rbv_graphs_clear(rbv_ctx* c, const void* map);   // drop captured graphs (all, or those that touch `map`)

static thread_local char g_err[512] = "";
void rbv_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* rbv_last_error(void) { return g_err; }

// ---- defaults (camera.hpp:26-37, edge_detector.hpp:19-32, edge_map.hpp:19-26, core.hpp:82-95) ----
extern "C" void rbv_default_camera(rbv_camera* cam) {
  const float fx = 458.654f, fy = 457.296f;
  cam->fm = (float)(0.5 * (double)(fx + fy));
  cam->cx = 367.215f;
  cam->cy = 248.375f;
  cam->rows = 480;
  cam->cols = 752;
}
extern "C" void rbv_default_detector_cfg(rbv_detector_cfg* c) {
  c->keylines_ref = 12000; c->keylines_max = 16000; c->pos_neg_threshold = 0.4f; c->dog_threshold = 0.095259868922420f;
  c->threshold = 0.01f; c->gain = 5e-7f; c->max_threshold = 0.5f; c->min_threshold = 0.005f;
}
extern "C" void rbv_default_map_cfg(rbv_map_cfg* c) {
  c->pixel_uncertainty_match = 2.0f; c->match_threshold_norm = 1.0f; c->match_threshold_angle = 45.0f; c->regularization_threshold = 0.5f;
}
extern "C" void rbv_default_core_cfg(rbv_core_cfg* c) {
  c->search_range = 40.0f; c->reweight_distance = 2.0f; c->match_treshold = 0.5f; c->min_match_threshold = 0; c->iterations = 5;
  c->global_min_matches_threshold = 500; c->pixel_uncertainty = 1.0f; c->quantile_cutoff = 0.9f; c->quantile_num_bins = 100;
  c->reshape_q_abs = 1e-4f;
}

// edge_detector.cpp:33-36
extern "C" float rbv_servo_threshold(const rbv_detector_cfg* cfg, float threshold, int last_keyline_count) {
  if (cfg->gain > 0) {
    threshold -= cfg->gain * (float)(cfg->keylines_ref - last_keyline_count);
    threshold = (threshold > cfg->max_threshold) ? cfg->max_threshold : ((threshold < cfg->min_threshold) ? cfg->min_threshold : threshold);
  }
  return threshold;
}

// Kovesi box widths, scale_space.cpp:20-35
static void kovesi_widths(float sigma, int n, int* widths, float* sigma_true) {
  const float w_ideal = (float)sqrt(12.0 * sigma * sigma / (float)(n + 1));
  int wl = (int)w_ideal;
  if ((wl / 2) * 2 == wl) --wl;
  const int m = (int)roundf(((float)(3 * n + 4 * n * wl + n * wl * wl) - 12 * sigma * sigma) / (float)(4 + 4 * wl));
  for (int i = 0; i < n; ++i) widths[i] = (i < m) ? wl : wl + 2;
  *sigma_true = (float)sqrt((m * wl * wl + (n - m) * (wl + 2.0) * (wl + 2.0) - n) / 12.0);
}

// 3x25 pseudo inverse of the 5x5 plane-fit design matrix, edge_detector.cpp:55-68
static void plane_fit_pinv(float* pinv75) {
  float phi[25][3], ptp[9], ipt[9];
  int k = 0;
  for (int r = -2; r <= 2; ++r)
    for (int q = -2; q <= 2; ++q, ++k) { phi[k][0] = (float)q; phi[k][1] = (float)r; phi[k][2] = 1.0f; }
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) {
      float acc = 0.0f;
      for (k = 0; k < 25; ++k) acc += phi[k][a] * phi[k][b];
      ptp[a * 3 + b] = acc;
    }
  rbv_inv3(ptp, ipt);
  for (int a = 0; a < 3; ++a)
    for (k = 0; k < 25; ++k) {
      float acc = 0.0f;
      for (int b = 0; b < 3; ++b) acc += ipt[a * 3 + b] * phi[k][b];
      pinv75[a * 25 + k] = acc;
    }
}

// ---- per-kernel launch accounting and CUDA-event timing ------------------------------------------------------
static const char* const g_kernel_names[KID_COUNT] = {
#define X(n) #n,
    RBV_KERNEL_LIST(X)
#undef X
};
#define PROF_PAIRS 4096
struct ProfState {
  cudaEvent_t ev[PROF_PAIRS][2];
  int kid[PROF_PAIRS];
  int used, created;
  double ms[KID_COUNT];
  unsigned long long n[KID_COUNT];
};
static void prof_drain(rbv_ctx* c) {
  ProfState* p = static_cast<ProfState*>(c->prof);
  if (!p || !p->used) return;
  cudaEventSynchronize(p->ev[p->used - 1][1]);
  for (int i = 0; i < p->used; ++i) {
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, p->ev[i][0], p->ev[i][1]) == cudaSuccess) { p->ms[p->kid[i]] += ms; ++p->n[p->kid[i]]; }
  }
  p->used = 0;
}
void rbv_prof_begin(rbv_ctx* c, int kid) {
  ++c->launches;
  ++c->klaunch[kid];
  if (c->prof_mode == -1 || (c->prof_mode >= 0 && c->prof_mode != kid)) return;
  ProfState* p = static_cast<ProfState*>(c->prof);
  if (p->used == PROF_PAIRS) prof_drain(c);
  if (p->used == p->created) {
    cudaEventCreate(&p->ev[p->created][0]);
    cudaEventCreate(&p->ev[p->created][1]);
    ++p->created;
  }
  p->kid[p->used] = kid;
  cudaEventRecord(p->ev[p->used][0], c->stream);
}
void rbv_prof_end(rbv_ctx* c, int kid) {
  if (c->prof_mode == -1 || (c->prof_mode >= 0 && c->prof_mode != kid)) return;
  ProfState* p = static_cast<ProfState*>(c->prof);
  cudaEventRecord(p->ev[p->used][1], c->stream);
  ++p->used;
}
extern "C" int rbv_kernel_count(void) { return KID_COUNT; }
extern "C" const char* rbv_kernel_name(int kid) { return (kid >= 0 && kid < KID_COUNT) ? g_kernel_names[kid] : ""; }
extern "C" int rbv_profile_set(rbv_ctx* c, int mode) {
  if (!c || mode < -2 || mode >= KID_COUNT) return RBV_ERR_ARG;
  LOCK_CTX(c);
  prof_drain(c);
  if (!c->prof) { ProfState* p = new ProfState(); memset(p, 0, sizeof(*p)); c->prof = p; }
  c->prof_mode = mode;
  return RBV_OK;
}
extern "C" int rbv_profile_reset(rbv_ctx* c) {
  if (!c) return RBV_ERR_ARG;
  LOCK_CTX(c);
  prof_drain(c);
  if (c->prof) { ProfState* p = static_cast<ProfState*>(c->prof); memset(p->ms, 0, sizeof(p->ms)); memset(p->n, 0, sizeof(p->n)); }
  memset(c->klaunch, 0, sizeof(c->klaunch));
  return RBV_OK;
}
extern "C" int rbv_profile_read(rbv_ctx* c, int kid, double* total_ms, unsigned long long* timed_launches, unsigned long long* launches) {
  if (!c || kid < 0 || kid >= KID_COUNT) return RBV_ERR_ARG;
  LOCK_CTX(c);
  prof_drain(c);
  ProfState* p = static_cast<ProfState*>(c->prof);
  if (total_ms) *total_ms = p ? p->ms[kid] : 0.0;
  if (timed_launches) *timed_launches = p ? p->n[kid] : 0ull;
  if (launches) *launches = c->klaunch[kid];
  return RBV_OK;
}

__global__ void k_fill(uint4* __restrict__ dst4, unsigned* __restrict__ tail, unsigned value, size_t n4, int ntail) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const uint4 v = make_uint4(value, value, value, value);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) dst4[i] = v;
  if (blockIdx.x == 0 && (int)threadIdx.x < ntail) tail[threadIdx.x] = value;
}
int rbv_fill32(rbv_ctx* c, void* dst, unsigned value, size_t n_words) {
  if (!n_words) return RBV_OK;
  if (((uintptr_t)dst & 15) != 0) { rbv_set_error("rbv_fill32: unaligned destination"); return RBV_ERR_ARG; }
  const size_t n4 = n_words / 4;
  const int ntail = (int)(n_words - n4 * 4);
  size_t blocks = (n4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  RBV_LAUNCH(c, k_fill, ((unsigned)blocks), (256), 0, (uint4*)dst, (unsigned*)dst + n4 * 4, value, n4, ntail);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

#define STAGE_SLOTS 16
struct Stage {  // ring of pinned staging slots for per-call lane inputs
  float* host[STAGE_SLOTS];
  cudaEvent_t ev[STAGE_SLOTS];
  int next;
};
static Stage* stage_of(rbv_ctx* c) { return reinterpret_cast<Stage*>(c->h_in); }

// Synchronous-looking copies IN STREAM ORDER on the context's stream. A plain cudaMemcpy from pageable memory may return
// while the DMA is still in flight and is not ordered against a non-blocking stream, so a kernel launched right after it
// on c->stream could read stale bytes.
static cudaError_t copy_sync(rbv_ctx* c, void* dst, const void* src, size_t bytes, cudaMemcpyKind kind) {
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, kind, c->stream);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(c->stream);
}

template <typename T>
static cudaError_t dalloc(T** p, size_t n) { return cudaMalloc((void**)p, n * sizeof(T)); }

// rbv_track_imu's own pinned result block (it drops the context lock while it waits for the device, so it must not share the
// staging buffers of the entry points another thread may call meanwhile): [B] LMState | [B][OUT_STRIDE] float | [B][CNT_STRIDE] int | [B] rbv_imu_out
static size_t trk_block_bytes(size_t B) { return B * (sizeof(LMState) + sizeof(float) * OUT_STRIDE + sizeof(int) * CNT_STRIDE + sizeof(rbv_imu_out)); }
static LMState* trk_lm(rbv_ctx* c) { return reinterpret_cast<LMState*>(c->h_trk); }
static float* trk_out(rbv_ctx* c) { return reinterpret_cast<float*>(trk_lm(c) + c->P.lanes); }
static int* trk_cnt(rbv_ctx* c) { return reinterpret_cast<int*>(trk_out(c) + (size_t)c->P.lanes * OUT_STRIDE); }
static rbv_imu_out* trk_imu(rbv_ctx* c) { return reinterpret_cast<rbv_imu_out*>(trk_cnt(c) + (size_t)c->P.lanes * CNT_STRIDE); }

static int create_body(rbv_ctx* c, const rbv_camera* cam, const rbv_detector_cfg* det, const rbv_map_cfg* map, const rbv_core_cfg* core,
                       int device, int lanes);
extern "C" int rbv_create(const rbv_camera* cam, const rbv_detector_cfg* det, const rbv_map_cfg* map, const rbv_core_cfg* core,
                          int device, int lanes, rbv_ctx** out) {
  if (!cam || !det || !map || !core || !out || lanes < 1 || lanes > 1024) { rbv_set_error("rbv_create: bad argument"); return RBV_ERR_ARG; }
  *out = nullptr;
  // pixel indices are int32 (rows * cols must fit with every lane offset held in size_t), rows / cols ride in 16-bit fields of the packed masks
  if (cam->rows < 32 || cam->cols < 32 || cam->rows > 32767 || cam->cols > 32767 || (long long)cam->rows * cam->cols > 0x7FFFFFFFll ||
      det->keylines_max < 1 || det->keylines_max >= (1 << 24)) {
    rbv_set_error("rbv_create: image size / keylines_max out of range"); return RBV_ERR_ARG;
  }
  if (core->iterations > RBV_MAX_LM_ITERS || core->search_range < 1 || core->search_range > 127 || core->quantile_num_bins < 1 ||
      core->quantile_num_bins > RBV_MAX_QBINS) { rbv_set_error("rbv_create: core config out of range"); return RBV_ERR_ARG; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { rbv_set_error("rbv_create: no CUDA device (there is no CPU fallback)"); return RBV_ERR_CUDA; }
  if (device < 0 || device >= ndev) { rbv_set_error("rbv_create: device %d out of range (%d devices)", device, ndev); return RBV_ERR_ARG; }
  cudaDeviceProp prop;
  RBV_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) { rbv_set_error("rbv_create: device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor); return RBV_ERR_CUDA; }
  RBV_CUDA(cudaSetDevice(device));
  rbv_ctx* c = new rbv_ctx();
  memset(c, 0, sizeof(*c));
  c->mtx = new std::recursive_mutex();
  c->device = device;
  const int rc = create_body(c, cam, det, map, core, device, lanes);
  if (rc != RBV_OK) { rbv_destroy(c); return rc; }   // a context is GBs of device and pinned memory: nothing of a failed one is left behind
  *out = c;
  return RBV_OK;
}

static int create_body(rbv_ctx* c, const rbv_camera* cam, const rbv_detector_cfg* det, const rbv_map_cfg* map, const rbv_core_cfg* core,
                       int device, int lanes) {
  c->cam = *cam; c->det = *det; c->mapc = *map; c->core = *core;
  c->device = device;
  c->reduce_mode = RBV_REDUCE_ORDERED;
  c->prof_mode = -1;
  float st[2];
  kovesi_widths(3.56359f, 3, c->widths[0], &st[0]);                       // scale_space.cpp:186
  kovesi_widths((float)(st[0] * 1.2599), 3, c->widths[1], &st[1]);
  for (int s = 0; s < 2; ++s) for (int p = 0; p < 3; ++p) if (c->widths[s][p] > 15) { rbv_set_error("box width > 15"); return RBV_ERR_ARG; }
  DevParams& P = c->P;
  P.fm = cam->fm; P.cx = cam->cx; P.cy = cam->cy; P.rows = cam->rows; P.cols = cam->cols; P.npix = cam->rows * cam->cols;
  P.keylines_max = det->keylines_max;
  P.kmax = (det->keylines_max + 31) & ~31;
  P.lanes = lanes;
  P.pn_threshold = (float)((2.0 * 2 + 1.0) * (2.0 * 2 + 1.0)) * det->pos_neg_threshold;   // edge_detector.cpp:69
  P.dog_threshold = det->dog_threshold;
  P.pixel_uncertainty_match = map->pixel_uncertainty_match; P.match_threshold_norm = map->match_threshold_norm;
  P.cang_min_edge = (float)cos(map->match_threshold_angle * M_PI / 180.0);                // edge_map.cpp:104
  P.regularization_threshold = map->regularization_threshold;
  P.search_range = core->search_range; P.reweight_distance = core->reweight_distance; P.match_treshold = core->match_treshold;
  P.pixel_uncertainty = core->pixel_uncertainty; P.quantile_cutoff = core->quantile_cutoff; P.reshape_q_abs = core->reshape_q_abs;
  P.quantile_num_bins = core->quantile_num_bins; P.iterations = (int)core->iterations;
  P.field_r0 = (int)-core->search_range;                                                   // core.hpp:49: for(int r = -search_range_; r < search_range_; ++r)
  P.field_nr = 0;
  for (int r = P.field_r0; r < core->search_range; ++r) ++P.field_nr;
  P.min_matches = 0;  // std::min(min_match_threshold, frame_count_) with frame_count_ never incremented (core.cpp:22,91)
  plane_fit_pinv(P.pinv);
  const size_t N = (size_t)P.npix, B = (size_t)lanes, K = (size_t)P.kmax;
  {
    // the context's own (tracker) stream outranks the front-end stream: the tracker is the longer chain of a frame, so when both
    // streams have blocks waiting (RBV_TUNE_FE_STREAM) its blocks go first (measured at 64 lanes: 53.6k -> 57.0k frames/s).
    // A caller's stream (rbv_set_stream) should be created with a higher-than-default priority for the same effect.
    int lo = 0, hi = 0;
    RBV_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    RBV_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, hi < lo ? lo - 1 : lo));
  }
  c->own_stream = true;
  { int rc = fe_init(c); if (rc != RBV_OK) return rc; }
  { int rc = tr_init(c); if (rc != RBV_OK) return rc; }
  const int wpr = (P.cols + 31) / 32;
  RBV_CUDA(dalloc(&c->img, B * N));
  RBV_CUDA(dalloc(&c->rowsum, 4 * B * N));   // rowsum [2][B][N] + bufB [2][B][N] contiguous (detect reuses it as float4 [B][N])
  c->bufB = c->rowsum + 2 * B * N;
  RBV_CUDA(dalloc(&c->bufA, 2 * B * N));
  RBV_CUDA(dalloc(&c->flags, B * P.rows * wpr));
  RBV_CUDA(dalloc(&c->rowcount, B * P.rows));
  RBV_CUDA(dalloc(&c->rowoffset, B * P.rows));
  RBV_CUDA(dalloc(&c->thr_dev, B));
  RBV_CUDA(dalloc(&c->last_count, B));
  {
    std::vector<float> t0(B, det->threshold);
    RBV_CUDA(cudaMemcpy(c->thr_dev, t0.data(), sizeof(float) * B, cudaMemcpyHostToDevice));
    RBV_CUDA(cudaMemset(c->last_count, 0, sizeof(int) * B));
  }
  RBV_CUDA(dalloc(&c->field, B * N));
  RBV_CUDA(dalloc(&c->tgt, 2 * B * K));
  {   // ORDERED addends: chunked layout of the chain kernels (rbv_tracker.cu, OrdLayout) or the 27 planes of the first form (RBV_TUNE_ORDERED = 1)
    const size_t chunked = tr_ordered_floats(lanes, P.kmax), planes = 27 * B * K;
    RBV_CUDA(dalloc(&c->prod, chunked > planes ? chunked : planes));
  }
  RBV_CUDA(dalloc(&c->fiabs, B * K));
  RBV_CUDA(dalloc(&c->fwd_head, B * K));
  RBV_CUDA(dalloc(&c->fwd_next, B * K));
  RBV_CUDA(dalloc(&c->reg_r, B * K));
  RBV_CUDA(dalloc(&c->reg_s, B * K));
  RBV_CUDA(dalloc(&c->lm, B));
  RBV_CUDA(dalloc(&c->lane_in, B * IN_STRIDE));
  RBV_CUDA(dalloc(&c->lane_out, B * OUT_STRIDE));
  RBV_CUDA(dalloc(&c->lane_cnt, B * CNT_STRIDE));
  RBV_CUDA(dalloc(&c->probes, B));
  c->nblk = (P.kmax + 255) / 256;
  RBV_CUDA(dalloc(&c->partials, B * c->nblk * 27));
  RBV_CUDA(dalloc(&c->ticket, B));
  RBV_CUDA(dalloc(&c->blkagg, 2 * B * c->nblk));
  RBV_CUDA(cudaMemset(c->ticket, 0, sizeof(unsigned) * B));
  RBV_CUDA(dalloc(&c->lm_done, B));
  RBV_CUDA(cudaMemset(c->lm_done, 0, sizeof(int) * B));
  RBV_CUDA(dalloc(&c->ord_flags, B * c->nblk));
  RBV_CUDA(cudaMemset(c->ord_flags, 0, sizeof(int) * B * c->nblk));
  RBV_CUDA(cudaMemset(c->blkagg, 0, sizeof(float) * 2 * B * c->nblk));
  RBV_CUDA(cudaMemset(c->rowcount, 0, sizeof(int) * B * P.rows));          // kept clean by k_detect_scan
  RBV_CUDA(cudaMemset(c->fwd_head, 0xFF, sizeof(int) * B * K));            // kept empty (-1) by k_fwd_resolve
  RBV_CUDA(cudaMemset(c->lm, 0, sizeof(LMState) * B));
  RBV_CUDA(cudaMemset(c->lane_in, 0, sizeof(float) * B * IN_STRIDE));
  RBV_CUDA(cudaMemset(c->lane_out, 0, sizeof(float) * B * OUT_STRIDE));
  RBV_CUDA(cudaMemset(c->lane_cnt, 0, sizeof(int) * B * CNT_STRIDE));
  RBV_CUDA(cudaMemset(c->probes, 0, sizeof(unsigned long long) * B));
  RBV_CUDA(cudaMemset(c->field, 0xFF, sizeof(unsigned) * B * N));
  Stage* stg = new Stage();
  for (int i = 0; i < STAGE_SLOTS; ++i) {
    RBV_CUDA(cudaMallocHost((void**)&stg->host[i], sizeof(float) * B * IN_STRIDE));
    RBV_CUDA(cudaEventCreateWithFlags(&stg->ev[i], cudaEventDisableTiming));
  }
  stg->next = 0;
  c->h_in = reinterpret_cast<float*>(stg);
  RBV_CUDA(cudaMallocHost((void**)&c->h_out, sizeof(float) * B * OUT_STRIDE));
  RBV_CUDA(cudaMallocHost((void**)&c->h_cnt, sizeof(int) * B * (CNT_STRIDE + 2)));
  RBV_CUDA(cudaMallocHost((void**)&c->h_lm, sizeof(LMState) * B));
  RBV_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
  {
    // stream priorities (0 = default = lowest, negative = higher): the tracker is the longer chain of a frame, so its second branch
    // may be given precedence over the front-end stream when both have blocks waiting (RBV_SIDE_PRIORITY / RBV_FE_PRIORITY)
    const char* sp = getenv("RBV_SIDE_PRIORITY");
    const char* fp = getenv("RBV_FE_PRIORITY");
    RBV_CUDA(cudaStreamCreateWithPriority(&c->side_stream, cudaStreamNonBlocking, sp ? atoi(sp) : 0));
    RBV_CUDA(cudaStreamCreateWithPriority(&c->fe_stream, cudaStreamNonBlocking, fp ? atoi(fp) : 0));
  }
  RBV_CUDA(cudaEventCreateWithFlags(&c->ev_tracker, cudaEventDisableTiming));
  RBV_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  RBV_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  RBV_CUDA(cudaEventCreateWithFlags(&c->ev_upload, cudaEventDisableTiming));
  for (int i = 0; i < RBV_RAW_SLOTS; ++i) {
    RBV_CUDA(cudaMalloc(&c->raw[i], sizeof(float) * B * N));
    RBV_CUDA(cudaEventCreateWithFlags(&c->raw_ready[i], cudaEventDisableTiming));
    RBV_CUDA(cudaEventCreateWithFlags(&c->raw_free[i], cudaEventDisableTiming));
  }
  RBV_CUDA(cudaMallocHost((void**)&c->h_misc, sizeof(int) * B * 2));
  RBV_CUDA(cudaMallocHost((void**)&c->h_trk, trk_block_bytes(B)));
  RBV_CUDA(cudaDeviceSynchronize());   // the memsets above ran on the legacy stream; the context's stream is non-blocking
  return RBV_OK;
}

extern "C" void rbv_destroy(rbv_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  void* dev[] = {c->img, c->rowsum, c->bufA, c->flags, c->rowcount, c->rowoffset, c->thr_dev, c->last_count, c->field, c->prod, c->fiabs,
                 c->fwd_head, c->fwd_next, c->reg_r, c->reg_s, c->lm, c->lane_in, c->lane_out, c->lane_cnt, c->probes,
                 c->partials, c->ticket, c->blkagg, c->tgt, c->ord_flags, c->lm_done};
  for (void* p : dev) cudaFree(p);
  Stage* stg = stage_of(c);
  if (stg) {
    for (int i = 0; i < STAGE_SLOTS; ++i) { cudaFreeHost(stg->host[i]); cudaEventDestroy(stg->ev[i]); }
    delete stg;
  }
  if (c->graphs) rbv_graphs_clear(c, nullptr);
  if (c->prof) {
    ProfState* p = static_cast<ProfState*>(c->prof);
    for (int i = 0; i < p->created; ++i) { cudaEventDestroy(p->ev[i][0]); cudaEventDestroy(p->ev[i][1]); }
    delete p;
  }
  cudaFreeHost(c->h_out); cudaFreeHost(c->h_cnt); cudaFreeHost(c->h_lm); cudaFreeHost(c->h_misc); cudaFreeHost(c->h_trk);
  if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
  if (c->side_stream) { cudaStreamSynchronize(c->side_stream); cudaStreamDestroy(c->side_stream); }
  if (c->fe_stream) { cudaStreamSynchronize(c->fe_stream); cudaStreamDestroy(c->fe_stream); }
  if (c->ev_tracker) cudaEventDestroy(c->ev_tracker);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->ev_upload) cudaEventDestroy(c->ev_upload);
  if (c->undist_map) cudaFree(c->undist_map);
  for (int i = 0; i < RBV_RAW_SLOTS; ++i) {
    cudaFree(c->raw[i]);
    if (c->rawf[i]) cudaFree(c->rawf[i]);
    if (c->raw_ready[i]) cudaEventDestroy(c->raw_ready[i]);
    if (c->raw_free[i]) cudaEventDestroy(c->raw_free[i]);
  }
  if (c->own_stream) cudaStreamDestroy(c->stream);
  cudaFree(c->imu_state); cudaFree(c->imu_in); cudaFree(c->imu_out); cudaFree(c->imu_stage); cudaFreeHost(c->imu_host);
  delete static_cast<std::recursive_mutex*>(c->mtx);
  delete c;
}

extern "C" int rbv_sync(rbv_ctx* c) {
  if (!c) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  if (c->fe_stream) RBV_CUDA(cudaStreamSynchronize(c->fe_stream));
  return RBV_OK;
}
extern "C" int rbv_set_reduce_mode(rbv_ctx* c, int mode) {
  if (!c || (mode != RBV_REDUCE_FAST && mode != RBV_REDUCE_ORDERED)) return RBV_ERR_ARG;
  LOCK_CTX(c);
  rbv_graphs_clear(c, nullptr);
  c->reduce_mode = mode;
  return RBV_OK;
}
extern "C" int rbv_set_tuning(rbv_ctx* c, int key, int value) {
  if (!c || key < 0 || key >= RBV_TUNE_COUNT || value < 0 || value > 64) return RBV_ERR_ARG;
  LOCK_CTX(c);
  cudaStreamSynchronize(c->stream);
  if (c->fe_stream) cudaStreamSynchronize(c->fe_stream);
  rbv_graphs_clear(c, nullptr);
  c->tune[key] = value;
  return RBV_OK;
}
extern "C" int rbv_lanes(const rbv_ctx* c) { return c ? c->P.lanes : 0; }
extern "C" int rbv_set_stream(rbv_ctx* c, void* cuda_stream) {
  if (!c) return RBV_ERR_ARG;
  LOCK_CTX(c);
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  rbv_graphs_clear(c, nullptr);
  if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
  c->stream = (cudaStream_t)cuda_stream;
  return RBV_OK;
}
extern "C" unsigned long long rbv_launch_count(const rbv_ctx* c) { return c ? c->launches : 0ull; }

// ---- maps ----
extern "C" int rbv_map_create(rbv_ctx* c, rbv_map** out) {
  if (!c || !out) return RBV_ERR_ARG;
  LOCK_CTX(c);
  RBV_CUDA(cudaSetDevice(c->device));
  const size_t B = c->P.lanes, K = c->P.kmax, N = c->P.npix;
  rbv_map* m = new rbv_map();
  memset(m, 0, sizeof(*m));
  // 5 float2 + 4 float + 5 int + 2 unsigned + 1 float (residual) per keyline, + mask + count + threshold
  const size_t per_kl = 5 * 8 + 4 * 4 + 5 * 4 + 2 * 4 + 4;
  m->slab_bytes = B * K * per_kl + B * N * 4 + B * 8 + 32 * 256;   // + slack for the 256-byte alignment of each array
  RBV_CUDA(cudaMalloc(&m->slab, m->slab_bytes));
  char* p = (char*)m->slab;
  auto take = [&](size_t bytes) { char* r = p; p += (bytes + 255) & ~(size_t)255; return r; };
  m->k.pos = (float2*)take(B * K * 8); m->k.pos_img = (float2*)take(B * K * 8); m->k.match_pos_img = (float2*)take(B * K * 8);
  m->k.grad = (float2*)take(B * K * 8); m->k.match_grad = (float2*)take(B * K * 8);
  m->k.gnorm = (float*)take(B * K * 4); m->k.match_gnorm = (float*)take(B * K * 4); m->k.rho = (float*)take(B * K * 4);
  m->k.srho = (float*)take(B * K * 4);
  m->k.id_prev = (int*)take(B * K * 4); m->k.id_next = (int*)take(B * K * 4); m->k.match_id = (int*)take(B * K * 4);
  m->k.match_fwd = (int*)take(B * K * 4); m->k.match_kf = (int*)take(B * K * 4);
  m->k.matches = (unsigned*)take(B * K * 4); m->k.pixel = (unsigned*)take(B * K * 4); m->k.residual = (float*)take(B * K * 4);
  m->mask = (int*)take(B * N * 4);
  m->count = (int*)take(B * 4);
  m->threshold = (float*)take(B * 4);
  RBV_CUDA(cudaMemsetAsync(m->slab, 0, m->slab_bytes, c->stream));
  RBV_CUDA(cudaMemsetAsync(m->mask, 0xFF, B * N * 4, c->stream));
  std::vector<float> thr(B, -1.0f);                                       // edge_map.cpp:17
  RBV_CUDA(cudaMemcpyAsync(m->threshold, thr.data(), B * 4, cudaMemcpyHostToDevice, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  m->ts_us = (uint64_t*)calloc(B, sizeof(uint64_t));
  RBV_CUDA(cudaEventCreateWithFlags(&m->ready, cudaEventDisableTiming));
  RBV_CUDA(cudaEventCreateWithFlags(&m->used, cudaEventDisableTiming));
  RBV_CUDA(cudaMallocHost((void**)&m->h_fe, B * 12));
  *out = m;
  return RBV_OK;
}

extern "C" void rbv_map_release(rbv_ctx* c, rbv_map* m) {
  if (!m) return;
  std::unique_lock<std::recursive_mutex> lk;
  if (c) lk = std::unique_lock<std::recursive_mutex>(*static_cast<std::recursive_mutex*>(c->mtx));
  if (c) {
    cudaSetDevice(c->device); cudaStreamSynchronize(c->stream);
    if (c->fe_stream) cudaStreamSynchronize(c->fe_stream);
    if (c->field_map == m) c->field_map = nullptr;
    rbv_graphs_clear(c, m);
  }
  if (m->field) cudaFree(m->field);
  if (m->tgt) cudaFree(m->tgt);
  if (m->ready) cudaEventDestroy(m->ready);
  if (m->used) cudaEventDestroy(m->used);
  if (m->h_fe) cudaFreeHost(m->h_fe);
  cudaFree(m->slab);
  free(m->ts_us);
  delete m;
}

// ---- per-call lane inputs: fill a pinned slot, async copy to lane_in ----
static float* stage_begin(rbv_ctx* c, int* slot) {
  Stage* s = stage_of(c);
  *slot = s->next;
  s->next = (s->next + 1) % STAGE_SLOTS;
  cudaEventSynchronize(s->ev[*slot]);
  memset(s->host[*slot], 0, sizeof(float) * (size_t)c->P.lanes * IN_STRIDE);
  return s->host[*slot];
}
static int stage_commit(rbv_ctx* c, int slot) {
  Stage* s = stage_of(c);
  RBV_CUDA(cudaMemcpyAsync(c->lane_in, s->host[slot], sizeof(float) * (size_t)c->P.lanes * IN_STRIDE, cudaMemcpyHostToDevice, c->stream));
  RBV_CUDA(cudaEventRecord(s->ev[slot], c->stream));
  return RBV_OK;
}
static void put(float* h, int lane, int slot, const float* src, int n) {
  if (src) memcpy(h + (size_t)lane * IN_STRIDE + slot, src + (size_t)lane * n, sizeof(float) * n);
}
static int fetch_out(rbv_ctx* c) {
  RBV_CUDA(cudaMemcpyAsync(c->h_out, c->lane_out, sizeof(float) * (size_t)c->P.lanes * OUT_STRIDE, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  return RBV_OK;
}
static int fetch_cnt(rbv_ctx* c) {
  RBV_CUDA(cudaMemcpyAsync(c->h_cnt, c->lane_cnt, sizeof(int) * (size_t)c->P.lanes * CNT_STRIDE, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  return RBV_OK;
}
// entry guard: null check, per-context lock (held to the end of the calling function; recursive because entry points call each
// other), device selection
#define CHECK_CTX(c)                                                        \
  if (!(c)) { rbv_set_error("null context"); return RBV_ERR_ARG; }          \
  LOCK_CTX(c);                                                              \
  RBV_CUDA(cudaSetDevice((c)->device))
#define TRY(x) do { int rc_ = (x); if (rc_ != RBV_OK) return rc_; } while (0)

// ---- front-end stream (SURVEY.md 8(b) "Threading": detect of frame n+1 next to the tracker of frames n-1, n) ---------------
// With RBV_TUNE_FE_STREAM = 1 the rbv_detect* calls enqueue on the context's own front-end stream instead of the tracker
// stream: the detector's buffers (img, rowsum, bufA/B, flags, row counts, thr_dev) are disjoint from the tracker's, a map
// carries an event recorded behind its detect, and every tracker / mirror entry waits for the events of the maps it is given.
// A detect without outputs then returns at once; rbv_detect_wait hands the keyline counts back later.
#define MAP_READY(c, m)                                                                                   \
  do {                                                                                                    \
    if ((m) && (m)->fe_pending) { RBV_CUDA(cudaStreamWaitEvent((c)->stream, (m)->ready, 0)); (m)->fe_pending = 0; } \
  } while (0)
// every tracker / mirror entry leaves an event on the tracker stream behind its work for the maps it was given (declared right
// behind MAP_READY, so it fires on every return path while the context lock is still held)
struct MapUse {
  rbv_ctx* c;
  rbv_map *a, *b;
  MapUse(rbv_ctx* ctx, rbv_map* m1, rbv_map* m2 = nullptr) : c(ctx), a(m1), b(m2) {}
  ~MapUse() {
    if (c->tune[RBV_TUNE_FE_STREAM] != 1) return;   // one stream: program order is enough
    if (a && a->used) { cudaEventRecord(a->used, c->stream); a->used_valid = 1; }
    if (b && b->used) { cudaEventRecord(b->used, c->stream); b->used_valid = 1; }
  }
};
struct FeScope {
  rbv_ctx* c;
  cudaStream_t saved;
  bool active;
  FeScope(rbv_ctx* ctx, rbv_map* target) : c(ctx), saved(ctx->stream), active(false) {
    if (c->tune[RBV_TUNE_FE_STREAM] == 1 && c->fe_stream && c->stream != c->fe_stream && c->prof_mode != -2) {
      // tracker work that was given the map this detect overwrites comes first (other tracker work may still be running: the
      // reference's estimation thread works on maps n-1, n while its acquisition thread fills map n+1)
      if (target && target->used_valid) cudaStreamWaitEvent(c->fe_stream, target->used, 0);
      c->stream = c->fe_stream;
      active = true;
    }
  }
  ~FeScope() { if (active) c->stream = saved; }
};


// ---- CUDA-graph cache ------------------------------------------------------------------------------------------------
// The kernel sequence of a stage group depends only on the context (geometry, tuning, reduce mode) and on which maps it
// works on: every per-call scalar travels through lane_in / thr_dev, every loop count is fixed (LM iterations), and host
// state that changes inside a group (tv_parity) returns to its start value. So the sequence is captured once per
// (group, map a, map b, flag) and replayed with ONE cudaGraphLaunch instead of 10-30 kernel launches; with one lane the
// frame is launch-latency bound and this is what the north star's "CUDA-graph-captured iterations" asks for.
// Not used while per-kernel event profiling is on (events cannot bracket kernels inside a graph) or with RBV_TUNE_GRAPH = 1.
struct GraphEntry {
  int kind;
  const void *a, *b;
  int flag;
  cudaGraphExec_t exec;
  unsigned long long launches;
  unsigned long long klaunch[KID_COUNT];
};
struct GraphCache {
  std::vector<GraphEntry> e;
};
void rbv_graphs_clear(rbv_ctx* c, const void* map) {
  GraphCache* g = static_cast<GraphCache*>(c->graphs);
  if (!g) return;
  for (size_t i = 0; i < g->e.size();) {
    if (!map || g->e[i].a == map || g->e[i].b == map) { cudaGraphExecDestroy(g->e[i].exec); g->e.erase(g->e.begin() + i); }
    else ++i;
  }
  if (!map) { delete g; c->graphs = nullptr; }
}
enum { GK_DETECT = 1, GK_TRACK1 = 2, GK_TRACK2 = 3 };
// scope of a fused tracker sequence (see rbv_ctx::seq_fused)
struct SeqFused {
  rbv_ctx* c;
  SeqFused(rbv_ctx* ctx, int min_matches, rbv_map* new_map = nullptr) : c(ctx) { c->seq_fused = 1; c->seq_min_matches = min_matches; c->seq_new = new_map; }
  ~SeqFused() { c->seq_fused = 0; c->seq_min_matches = 0; c->seq_new = nullptr; c->seq_reg_pending = 0; c->seq_resolve_agg = nullptr; }
};
template <class Body>
static int run_graphed(rbv_ctx* c, int kind, const void* a, const void* b, int flag, Body body) {
  if (c->tune[RBV_TUNE_GRAPH] == 1 || c->prof_mode != -1) return body();
  if (!c->graphs) c->graphs = new GraphCache();
  GraphCache* g = static_cast<GraphCache*>(c->graphs);
  for (GraphEntry& e : g->e)
    if (e.kind == kind && e.a == a && e.b == b && e.flag == flag) {
      c->launches += e.launches;
      for (int k = 0; k < KID_COUNT; ++k) c->klaunch[k] += e.klaunch[k];
      RBV_CUDA(cudaGraphLaunch(e.exec, c->stream));
      return RBV_OK;
    }
  GraphEntry e;
  memset(&e, 0, sizeof(e));
  e.kind = kind; e.a = a; e.b = b; e.flag = flag;
  const unsigned long long l0 = c->launches;
  unsigned long long k0[KID_COUNT];
  for (int k = 0; k < KID_COUNT; ++k) k0[k] = c->klaunch[k];
  RBV_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed));
  const int rc = body();
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
  if (rc != RBV_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
  if (ce != cudaSuccess || !graph) { rbv_set_error("graph capture failed: %s", cudaGetErrorString(ce)); return RBV_ERR_CUDA; }
  const cudaError_t ie = cudaGraphInstantiate(&e.exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ie != cudaSuccess) { rbv_set_error("cudaGraphInstantiate: %s", cudaGetErrorString(ie)); return RBV_ERR_CUDA; }
  e.launches = c->launches - l0;
  for (int k = 0; k < KID_COUNT; ++k) e.klaunch[k] = c->klaunch[k] - k0[k];
  if (g->e.size() >= 16) { cudaGraphExecDestroy(g->e.front().exec); g->e.erase(g->e.begin()); }
  g->e.push_back(e);
  RBV_CUDA(cudaGraphLaunch(e.exec, c->stream));
  return RBV_OK;
}

// ---- distance field in the front end (RBV_TUNE_FIELD_IN_FE) --------------------------------------------------------------------
// DistanceField::build (core.hpp:37-59) reads only what the detector wrote (pos, gradient, gradient_norm, the auto threshold): none of
// it changes before the tracker's first half uses the field, so the build can follow the detect on the front-end stream instead of
// heading the tracker chain - the longest tracker kernel leaves the critical path. The field then belongs to the map (the front end
// may be frames ahead of the tracker); the tracker's kernels are pointed at it for the duration of the enqueue.
static int map_field_alloc(rbv_ctx* c, rbv_map* m) {
  if (m->field && m->tgt) return RBV_OK;
  const size_t B = c->P.lanes, K = c->P.kmax, N = c->P.npix;
  if (!m->field) { RBV_CUDA(cudaMalloc((void**)&m->field, sizeof(unsigned) * B * N)); RBV_CUDA(cudaMemset(m->field, 0xFF, sizeof(unsigned) * B * N)); }
  if (!m->tgt) RBV_CUDA(cudaMalloc((void**)&m->tgt, sizeof(float4) * 2 * B * K));
  return RBV_OK;
}
struct FieldScope {   // point the context's field / target slots at a map's own buffers
  rbv_ctx* c;
  unsigned* f0;
  float4* t0;
  FieldScope(rbv_ctx* ctx, rbv_map* m, bool on) : c(ctx), f0(ctx->field), t0(ctx->tgt) { if (on) { c->field = m->field; c->tgt = m->tgt; } }
  ~FieldScope() { c->field = f0; c->tgt = t0; }
};

// edge_detector.cpp:33-36 on the device, for callers that pass threshold = NULL: config_->threshold of every lane lives in thr_dev,
// keylines_count_ in last_count (left there by the previous detect). Same float32 statement as rbv_servo_threshold (-fmad=false).
// With the servo on the device a detect depends on nothing the host has to read back, so the front end can be enqueued frames ahead.
__global__ void k_servo(float* __restrict__ thr, const int* __restrict__ last_count, int B, float gain, int keylines_ref, float tmin, float tmax) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float t = thr[b];
  if (gain > 0) {
    t -= gain * (float)(keylines_ref - last_count[b]);
    t = (t > tmax) ? tmax : ((t < tmin) ? tmin : t);
  }
  thr[b] = t;
}

// ---- front end ----
extern "C" int rbv_detect_device(rbv_ctx* c, rbv_map* map, const float* d_images, const uint64_t* ts_us, const float* threshold,
                                 int* n_keylines, float* auto_threshold) {
  CHECK_CTX(c);
  if (!map || !d_images) { rbv_set_error("rbv_detect: null argument"); return RBV_ERR_ARG; }
  FeScope fe(c, map);
  const bool on_fe = c->stream == c->fe_stream;
  const int B = c->P.lanes;
  if (threshold) {
    int slot;
    float* h = stage_begin(c, &slot);   // reuse a staging slot for the B thresholds
    memcpy(h, threshold, sizeof(float) * B);
    Stage* s = stage_of(c);
    RBV_CUDA(cudaMemcpyAsync(c->thr_dev, h, sizeof(float) * B, cudaMemcpyHostToDevice, c->stream));
    RBV_CUDA(cudaEventRecord(s->ev[slot], c->stream));
  } else {                              // threshold servo on the device from the previous detect's keyline counts
    RBV_LAUNCH(c, k_servo, ((unsigned)((B + 127) / 128)), (128), 0, c->thr_dev, c->last_count, B, c->det.gain, c->det.keylines_ref,
               c->det.min_threshold, c->det.max_threshold);
    RBV_CUDA(cudaGetLastError());
  }
  if (ts_us) memcpy(map->ts_us, ts_us, sizeof(uint64_t) * B);
  // the one kernel that reads the caller's frame buffer is launched directly; everything behind it works on the
  // context's own buffers and is replayed from a graph
  // the distance field of this map follows its detect on the front-end stream: k_detect_emit then also writes the packed target records
  // (one launch less; same values as k_pack_targets computes from what emit stores)
  const bool fe_field = on_fe && c->tune[RBV_TUNE_FIELD_IN_FE] == 1;
  if (fe_field) TRY(map_field_alloc(c, map));
  TRY(fe_scale_space(c, d_images, 1));
  TRY(run_graphed(c, GK_DETECT, map, nullptr, fe_uses_fused_box(c, d_images) + (fe_field ? 8 : 0), [&]() -> int {
    TRY(fe_scale_space(c, d_images, 2));
    return fe_detect(c, map, fe_field ? map->tgt : nullptr);
  }));
  map->fe_valid = 0;
  map->field_built = 0;
  if (fe_field) {
    rbv_map* fm0 = c->field_map;
    int rc;
    { FieldScope fs(c, map, true); rc = tr_build_field(c, map, true); }
    c->field_map = fm0;   // host-side notion of the tracker's current field map: unchanged by a front end running ahead
    TRY(rc);
    map->field_built = 1;
  }
  RBV_CUDA(cudaMemcpyAsync(c->last_count, map->count, sizeof(int) * B, cudaMemcpyDeviceToDevice, c->stream));   // keylines_count_ for k_servo
  if (on_fe) {   // counts, auto thresholds and the thresholds used leave on the front-end stream (rbv_detect_wait), the map's event follows them
    RBV_CUDA(cudaMemcpyAsync(map->h_fe, map->count, sizeof(int) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaMemcpyAsync(map->h_fe + B, map->threshold, sizeof(float) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaMemcpyAsync(map->h_fe + 2 * B, c->thr_dev, sizeof(float) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaEventRecord(map->ready, c->stream));
    map->fe_pending = 1;
    map->fe_valid = 1;
  }
  if (n_keylines || auto_threshold) {
    int* hc = c->h_misc;     // not h_cnt: rbv_track_imu on another thread may be waiting for its counters (it drops the lock meanwhile)
    float* ht = reinterpret_cast<float*>(c->h_misc + B);
    RBV_CUDA(cudaMemcpyAsync(hc, map->count, sizeof(int) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaMemcpyAsync(ht, map->threshold, sizeof(float) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaStreamSynchronize(c->stream));
    if (n_keylines) memcpy(n_keylines, hc, sizeof(int) * B);
    if (auto_threshold) memcpy(auto_threshold, ht, sizeof(float) * B);
  }
  return RBV_OK;
}

// u8 -> float32 x 3.0 (Rebvio::imageCallback, rebvio.cpp:43: convertTo(CV_32FC1, 3.0)); exact in float32
__global__ void k_u8_to_f32(const uint4* __restrict__ src16, float4* __restrict__ dst, const unsigned char* __restrict__ src, float* __restrict__ dstf,
                            size_t n16, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    const uint4 v = src16[i];
    const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k)
      dst[i * 4 + k] = make_float4((float)(w[k] & 0xFFu) * 3.0f, (float)((w[k] >> 8) & 0xFFu) * 3.0f, (float)((w[k] >> 16) & 0xFFu) * 3.0f,
                                   (float)(w[k] >> 24) * 3.0f);
  }
  for (size_t i = n16 * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dstf[i] = (float)src[i] * 3.0f;
}

// ---- Camera::undistort on the device ------------------------------------------------------------
// Rebvio::imageCallback (rebvio.cpp:43-45) converts the camera frame to float32 x 3 and runs cv::undistort(K, D) on it on the
// host, holding the image mutex. cv::undistort = per-pixel map in double (initUndistortRectifyMap: normalise with inv(K), radial
// (k1, k2, k3) and tangential (p1, p2) distortion, project with K), quantised to 1/32 pixel (CV_16SC2 + CV_16UC1 maps,
// INTER_BITS = 5), then remap INTER_LINEAR / BORDER_CONSTANT 0 with the float32 weight table. The map depends on the camera
// only, so it is built once (k_undistort_map, 8 bytes per pixel: source x, y as int16, the four weights' table index) and every
// frame is one gather kernel that also does the x 3 conversion. oracle/undistort.py restates the same algorithm and is pinned
// bit for bit against cv2.undistort (tests/golden/undistort_euroc.npz).
struct UndistortK {
  double ifx, ifx_cx, ify, ify_cy, fx, fy, cx, cy, k1, k2, p1, p2, k3;
};
__global__ void k_undistort_map(uint2* __restrict__ map, int rows, int cols, const __grid_constant__ UndistortK U) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= cols) return;
  const double x = (double)j * U.ifx + U.ifx_cx, y = (double)i * U.ify + U.ify_cy;       // inv(K) * (j, i, 1)
  const double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2.0 * x * y;
  const double kr = 1.0 + ((U.k3 * r2 + U.k2) * r2 + U.k1) * r2;
  const double xd = x * kr + U.p1 * _2xy + U.p2 * (r2 + 2.0 * x2);
  const double yd = y * kr + U.p1 * (r2 + 2.0 * y2) + U.p2 * _2xy;
  const double u = U.fx * xd + U.cx, v = U.fy * yd + U.cy;
  const long long iu = __double2ll_rn(u * 32.0), iv = __double2ll_rn(v * 32.0);         // cvRound: half to even
  long long sx = iu >> 5, sy = iv >> 5;
  // anything that cannot touch the image is parked at one fully-outside position (result 0, as BORDER_CONSTANT gives)
  if (sx < -1 || sx > cols || sy < -1 || sy > rows) { sx = -2; sy = -2; }
  const unsigned wx = (unsigned)(iu & 31), wy = (unsigned)(iv & 31);
  map[(size_t)i * cols + j] = make_uint2(((unsigned)(sx & 0xFFFF)) | ((unsigned)(sy & 0xFFFF) << 16), wy * 32u + wx);
}
template <typename T> __device__ __forceinline__ float ud_fetch(const T* __restrict__ src, int rows, int cols, int yy, int xx) {
  if ((unsigned)yy >= (unsigned)rows || (unsigned)xx >= (unsigned)cols) return 0.0f;    // BORDER_CONSTANT, value 0
  return (float)src[(size_t)yy * cols + xx] * (sizeof(T) == 1 ? 3.0f : 1.0f);            // convertTo(CV_32F, 3.0) for camera bytes
}
__global__ void __launch_bounds__(256) k_undistort(const unsigned char* __restrict__ src8, const float* __restrict__ src32, float* __restrict__ dst,
                                                   const uint2* __restrict__ map, int rows, int cols) {
  const size_t N = (size_t)rows * cols;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if ((size_t)p >= N) return;
  const uint2 m = map[p];
  const int sx = (int)(short)(m.x & 0xFFFFu), sy = (int)(short)(m.x >> 16);
  const float fx = (float)(m.y & 31u) * 0.03125f, fy = (float)(m.y >> 5) * 0.03125f;
  const float cx0 = 1.0f - fx, cy0 = 1.0f - fy;
  const float w00 = cy0 * cx0, w01 = cy0 * fx, w10 = fy * cx0, w11 = fy * fx;
  float s00, s01, s10, s11;
  if (src8) {
    const unsigned char* s = src8 + (size_t)blockIdx.y * N;
    s00 = ud_fetch(s, rows, cols, sy, sx); s01 = ud_fetch(s, rows, cols, sy, sx + 1);
    s10 = ud_fetch(s, rows, cols, sy + 1, sx); s11 = ud_fetch(s, rows, cols, sy + 1, sx + 1);
  } else {
    const float* s = src32 + (size_t)blockIdx.y * N;
    s00 = ud_fetch(s, rows, cols, sy, sx); s01 = ud_fetch(s, rows, cols, sy, sx + 1);
    s10 = ud_fetch(s, rows, cols, sy + 1, sx); s11 = ud_fetch(s, rows, cols, sy + 1, sx + 1);
  }
  dst[(size_t)blockIdx.y * N + p] = ((s00 * w00 + s01 * w01) + s10 * w10) + s11 * w11;
}

extern "C" int rbv_set_undistort(rbv_ctx* c, const float* K4, const float* D5) {
  CHECK_CTX(c);
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  rbv_graphs_clear(c, nullptr);
  if (!K4) {                              // frames arrive undistorted again
    if (c->undist_map) { cudaFree(c->undist_map); c->undist_map = nullptr; }
    return RBV_OK;
  }
  if (!D5) { rbv_set_error("rbv_set_undistort: distortion coefficients missing"); return RBV_ERR_ARG; }
  if (c->P.rows > 32000 || c->P.cols > 32000) { rbv_set_error("rbv_set_undistort: image too large for 16-bit map coordinates"); return RBV_ERR_ARG; }
  if (!(K4[0] > 0.0f) || !(K4[1] > 0.0f)) { rbv_set_error("rbv_set_undistort: focal lengths must be positive"); return RBV_ERR_ARG; }
  UndistortK U;
  U.fx = K4[0]; U.fy = K4[1]; U.cx = K4[2]; U.cy = K4[3];                     // float32 K_, D_ widened to double (camera.hpp:39-40)
  U.ifx = 1.0 / U.fx; U.ify = 1.0 / U.fy; U.ifx_cx = -U.cx * U.ifx; U.ify_cy = -U.cy * U.ify;
  U.k1 = D5[0]; U.k2 = D5[1]; U.p1 = D5[2]; U.p2 = D5[3]; U.k3 = D5[4];
  if (!c->undist_map) RBV_CUDA(cudaMalloc(&c->undist_map, sizeof(uint2) * (size_t)c->P.npix));
  RBV_LAUNCH(c, k_undistort_map, (dim3((c->P.cols + 127) / 128, c->P.rows)), (128), 0, c->undist_map, c->P.rows, c->P.cols, U);
  RBV_CUDA(cudaGetLastError());
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  return RBV_OK;
}

// H2D of B frames (element size esz) into a dense [B][N] device buffer on stream st: one copy when the host frames are
// dense and back to back, else one copy per lane
static int upload_frames(rbv_ctx* c, void* dst, const void* const* images, size_t pitch_bytes, size_t esz, cudaStream_t st) {
  const size_t N = c->P.npix, row_bytes = esz * c->P.cols;
  const int B = c->P.lanes;
  if (pitch_bytes == 0) pitch_bytes = row_bytes;
  if (pitch_bytes < row_bytes) { rbv_set_error("frame upload: pitch smaller than a row"); return RBV_ERR_ARG; }
  bool dense = pitch_bytes == row_bytes;
  for (int b = 0; b < B; ++b) {
    if (!images[b]) { rbv_set_error("frame upload: null image for lane %d", b); return RBV_ERR_ARG; }
    if ((const char*)images[b] != (const char*)images[0] + (size_t)b * N * esz) dense = false;
  }
  if (dense) {
    RBV_CUDA(cudaMemcpyAsync(dst, images[0], (size_t)B * N * esz, cudaMemcpyHostToDevice, st));
    return RBV_OK;
  }
  for (int b = 0; b < B; ++b) {
    char* d = (char*)dst + (size_t)b * N * esz;
    if (pitch_bytes == row_bytes) RBV_CUDA(cudaMemcpyAsync(d, images[b], N * esz, cudaMemcpyHostToDevice, st));
    else RBV_CUDA(cudaMemcpy2DAsync(d, row_bytes, images[b], pitch_bytes, row_bytes, c->P.rows, cudaMemcpyHostToDevice, st));
  }
  return RBV_OK;
}

extern "C" int rbv_prefetch_frames(rbv_ctx* c, const void* const* images, size_t pitch_bytes, int is_u8) {
  CHECK_CTX(c);
  if (!images) { rbv_set_error("rbv_prefetch_frames: null images"); return RBV_ERR_ARG; }
  const int slot = c->raw_next;
  RBV_CUDA(cudaStreamWaitEvent(c->copy_stream, c->raw_free[slot], 0));   // the front end is done with what was in this slot
  TRY(upload_frames(c, c->raw[slot], images, pitch_bytes, is_u8 ? 1 : 4, c->copy_stream));
  c->raw_conv[slot] = 0;
  if (is_u8 && !c->undist_map) {   // convertTo(CV_32FC1, 3.0) (rebvio.cpp:43) behind the upload, on the copy stream: one kernel less on the front-end chain
    const size_t n = (size_t)c->P.lanes * c->P.npix;
    if (!c->rawf[slot]) RBV_CUDA(cudaMalloc((void**)&c->rawf[slot], sizeof(float) * n));
    cudaStream_t s0 = c->stream;
    c->stream = c->copy_stream;
    RBV_LAUNCH(c, k_u8_to_f32, (148 * 8), (256), 0, (const uint4*)c->raw[slot], (float4*)c->rawf[slot], (const unsigned char*)c->raw[slot],
               c->rawf[slot], n / 16, n);
    c->stream = s0;
    RBV_CUDA(cudaGetLastError());
    c->raw_conv[slot] = 1;
  }
  RBV_CUDA(cudaEventRecord(c->raw_ready[slot], c->copy_stream));
  c->raw_key[slot] = images[0];
  c->raw_u8[slot] = is_u8 ? 1 : 0;
  c->raw_next = (slot + 1) % RBV_RAW_SLOTS;
  return RBV_OK;
}

static int detect_host(rbv_ctx* c, rbv_map* map, const void* const* images, size_t pitch_bytes, int is_u8, const uint64_t* ts_us,
                       const float* threshold, int* n_keylines, float* auto_threshold) {
  if (!images || !images[0]) { rbv_set_error("rbv_detect: null images"); return RBV_ERR_ARG; }
  int slot = -1;
  for (int i = 0; i < RBV_RAW_SLOTS; ++i)
    if (c->raw_key[i] == images[0] && c->raw_u8[i] == (is_u8 ? 1 : 0)) slot = i;
  const float* d_img;
  const bool converted = slot >= 0 && is_u8 && c->raw_conv[slot] && !c->undist_map;   // prefetched and already float32 x 3
  cudaEvent_t upload_ev = nullptr;       // fires when the caller's host frames have been read
  if (slot >= 0) {                       // staged by rbv_prefetch_frames: wait for the copy stream, no second copy
    RBV_CUDA(cudaStreamWaitEvent(c->stream, c->raw_ready[slot], 0));
    c->raw_key[slot] = nullptr;
    upload_ev = c->raw_ready[slot];
  } else if (is_u8) {                    // upload now, in stream order
    slot = c->raw_next;
    c->raw_next = (slot + 1) % RBV_RAW_SLOTS;
    c->raw_key[slot] = nullptr;
    RBV_CUDA(cudaStreamWaitEvent(c->stream, c->raw_ready[slot], 0));   // a stale prefetch into this slot must have landed first
    TRY(upload_frames(c, c->raw[slot], images, pitch_bytes, 1, c->stream));
    RBV_CUDA(cudaEventRecord(c->ev_upload, c->stream)); upload_ev = c->ev_upload;
  } else if (c->undist_map) {            // float frames that still need undistorting: stage them, the remap writes c->img
    slot = c->raw_next;
    c->raw_next = (slot + 1) % RBV_RAW_SLOTS;
    c->raw_key[slot] = nullptr;
    RBV_CUDA(cudaStreamWaitEvent(c->stream, c->raw_ready[slot], 0));
    TRY(upload_frames(c, c->raw[slot], images, pitch_bytes, 4, c->stream));
    RBV_CUDA(cudaEventRecord(c->ev_upload, c->stream)); upload_ev = c->ev_upload;
  } else {
    TRY(upload_frames(c, c->img, images, pitch_bytes, 4, c->stream));
    RBV_CUDA(cudaEventRecord(c->ev_upload, c->stream)); upload_ev = c->ev_upload;
  }
  if (c->undist_map) {                   // Camera::undistort (+ the x3 conversion for camera bytes) in one gather pass
    RBV_LAUNCH(c, k_undistort, (dim3((unsigned)((c->P.npix + 255) / 256), c->P.lanes)), (256), 0,
               is_u8 ? (const unsigned char*)c->raw[slot] : nullptr, is_u8 ? nullptr : (const float*)c->raw[slot], c->img, c->undist_map,
               c->P.rows, c->P.cols);
    RBV_CUDA(cudaGetLastError());
    d_img = c->img;
  } else if (converted) {
    d_img = c->rawf[slot];
  } else if (is_u8) {
    const size_t n = (size_t)c->P.lanes * c->P.npix;
    RBV_LAUNCH(c, k_u8_to_f32, (148 * 8), (256), 0, (const uint4*)c->raw[slot], (float4*)c->img, (const unsigned char*)c->raw[slot], c->img,
               n / 16, n);
    RBV_CUDA(cudaGetLastError());
    d_img = c->img;
  } else {
    d_img = slot >= 0 ? (const float*)c->raw[slot] : c->img;
  }
  const int rc = rbv_detect_device(c, map, d_img, ts_us, threshold, n_keylines, auto_threshold);
  if (slot >= 0) RBV_CUDA(cudaEventRecord(c->raw_free[slot], c->stream));
  // the header's promise: the caller's frames may be reused once this call returns. A copy from pinned memory is a real DMA that may still
  // be reading them, so the call waits for the upload (not for the kernels behind it); a prefetched upload was started a frame ago
  if (upload_ev) RBV_CUDA(cudaEventSynchronize(upload_ev));
  return rc;
}

extern "C" int rbv_detect(rbv_ctx* c, rbv_map* map, const float* const* images, size_t pitch_bytes, const uint64_t* ts_us,
                          const float* threshold, int* n_keylines, float* auto_threshold) {
  CHECK_CTX(c);
  FeScope fe(c, map);
  return detect_host(c, map, (const void* const*)images, pitch_bytes, 0, ts_us, threshold, n_keylines, auto_threshold);
}

extern "C" int rbv_detect_u8(rbv_ctx* c, rbv_map* map, const uint8_t* const* images, size_t pitch_bytes, const uint64_t* ts_us,
                             const float* threshold, int* n_keylines, float* auto_threshold) {
  CHECK_CTX(c);
  FeScope fe(c, map);
  return detect_host(c, map, (const void* const*)images, pitch_bytes, 1, ts_us, threshold, n_keylines, auto_threshold);
}

// keyline counts and auto thresholds of the most recent detect into `map` (both optional, length B). After a detect on the
// front-end stream this waits for that detect alone (the tracker stream keeps running); otherwise it reads the map.
extern "C" int rbv_detect_wait(rbv_ctx* c, rbv_map* map, int* n_keylines, float* auto_threshold, float* threshold_used) {
  if (!c) { rbv_set_error("null context"); return RBV_ERR_ARG; }
  std::unique_lock<std::recursive_mutex> ctx_lock_(*static_cast<std::recursive_mutex*>(c->mtx));   // outermost entry, see rbv_track_imu
  RBV_CUDA(cudaSetDevice(c->device));
  if (!map) return RBV_ERR_ARG;
  const size_t B = c->P.lanes;
  if (map->fe_valid) {
    {   // the wait itself needs nothing of the context: the estimation thread may enqueue its tracker call meanwhile
      cudaEvent_t ev = map->ready;
      ctx_lock_.unlock();
      const cudaError_t se = cudaEventSynchronize(ev);
      ctx_lock_.lock();
      cudaSetDevice(c->device);
      RBV_CUDA(se);
    }
    if (n_keylines) memcpy(n_keylines, map->h_fe, sizeof(int) * B);
    if (auto_threshold) memcpy(auto_threshold, map->h_fe + B, sizeof(float) * B);
    if (threshold_used) memcpy(threshold_used, map->h_fe + 2 * B, sizeof(float) * B);
    return RBV_OK;
  }
  if (n_keylines) TRY(rbv_map_size(c, map, n_keylines));
  if (auto_threshold) TRY(rbv_map_threshold(c, map, auto_threshold));
  if (threshold_used) RBV_CUDA(copy_sync(c, threshold_used, c->thr_dev, sizeof(float) * B, cudaMemcpyDeviceToHost));   // of the context's latest detect
  return RBV_OK;
}

// ---- tracker stages ----
extern "C" int rbv_build_distance_field(rbv_ctx* c, rbv_map* map) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map) return RBV_ERR_ARG;
  return tr_build_field(c, map);
}

extern "C" int rbv_rotate_keylines(rbv_ctx* c, rbv_map* map, const float* R9) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map || !R9) return RBV_ERR_ARG;
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_R, R9, 9);
  TRY(stage_commit(c, slot));
  return tr_rotate(c, map);
}

extern "C" int rbv_estimate_quantile(rbv_ctx* c, rbv_map* map, float percentile, int num_bins, float* out) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map) return RBV_ERR_ARG;
  TRY(tr_quantile(c, map, percentile, num_bins, 0));
  if (out) {
    TRY(fetch_out(c));
    for (int b = 0; b < c->P.lanes; ++b) out[b] = c->h_out[(size_t)b * OUT_STRIDE + OUT_Q];
  }
  return RBV_OK;
}

extern "C" int rbv_reset_residuals(rbv_ctx* c, rbv_map* map) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map) return RBV_ERR_ARG;
  return rbv_fill32(c, map->k.residual, 0u, (size_t)c->P.lanes * c->P.kmax);   // core.cpp:158
}

extern "C" int rbv_try_vel(rbv_ctx* c, rbv_map* map, const float* vel3, const float* sigma_rho_min, float* JtJ9, float* JtF3, float* score) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map || !vel3 || !sigma_rho_min) return RBV_ERR_ARG;
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) { put(h, b, IN_VEL, vel3, 3); put(h, b, IN_SRMIN, sigma_rho_min, 1); }
  TRY(stage_commit(c, slot));
  TRY(tr_try_vel(c, map, -1));
  if (JtJ9 || JtF3 || score) {
    TRY(fetch_out(c));
    for (int b = 0; b < c->P.lanes; ++b) {
      const float* o = c->h_out + (size_t)b * OUT_STRIDE;
      if (JtJ9) memcpy(JtJ9 + b * 9, o + OUT_JTJ, 36);
      if (JtF3) memcpy(JtF3 + b * 3, o + OUT_JTF, 12);
      if (score) score[b] = o[OUT_SCORE];
    }
  }
  return RBV_OK;
}

__global__ void k_lm_init(LMState* lm, const float* lane_in, int lanes) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= lanes) return;
  for (int i = 0; i < 3; ++i) lm[b].vel[i] = lane_in[(size_t)b * IN_STRIDE + IN_VEL + i];
}

// enqueue Core::minimizeVel (core.cpp:150-189) on the stream: quantile, residual reset, (1 + iterations) tryVel + LM steps
// part: 0 = everything, 1 = only what precedes the first tryVel (quantile, LM seed), 2 = only the tryVel / LM calls
static int enqueue_minimize(rbv_ctx* c, rbv_map* old_map, int part = 0) {
  // FAST mode with several keylines per thread: the quantile kernel also seeds the LM state, and call 0 of tryVel takes the
  // residuals as zero and writes every entry, so neither k_lm_init nor the residual fill is launched
  const bool lean = !tr_ordered_legacy(c) && (c->reduce_mode == RBV_REDUCE_ORDERED || c->tune[RBV_TUNE_TRYVEL] != 1);
  if (part != 2) {
    TRY(tr_quantile(c, old_map, c->P.quantile_cutoff, c->P.quantile_num_bins, lean ? 2 : 1));
    if (!lean) {
      TRY(rbv_fill32(c, old_map->k.residual, 0u, (size_t)c->P.lanes * c->P.kmax));
      RBV_LAUNCH(c, k_lm_init, ((c->P.lanes + 63) / 64), (64), 0, c->lm, c->lane_in, c->P.lanes);
    }
  }
  if (part != 1) {
    if (lean && tr_minimize_persist_wanted(c)) TRY(tr_minimize_persist(c, old_map));   // RBV_TUNE_ORDERED = 4: all calls in one launch
    else for (int t = 0; t <= c->P.iterations; ++t) TRY(tr_try_vel(c, old_map, t));
  }
  return RBV_OK;
}
static int fetch_lm(rbv_ctx* c, float* vel3, float* Rvel9, float* score) {
  RBV_CUDA(cudaMemcpyAsync(c->h_lm, c->lm, sizeof(LMState) * (size_t)c->P.lanes, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  const LMState* L = reinterpret_cast<const LMState*>(c->h_lm);
  for (int b = 0; b < c->P.lanes; ++b) {
    if (vel3) memcpy(vel3 + b * 3, L[b].vel, 12);
    if (Rvel9) memcpy(Rvel9 + b * 9, L[b].Rvel, 36);
    if (score) score[b] = L[b].F;
  }
  return RBV_OK;
}

extern "C" int rbv_minimize_vel(rbv_ctx* c, rbv_map* old_map, float* vel3, float* Rvel9, float* score) {
  CHECK_CTX(c);
  MAP_READY(c, old_map);
  MapUse map_use_(c, old_map);
  if (!old_map || !vel3) return RBV_ERR_ARG;
  if (!c->field_map) { rbv_set_error("minimizeVel before buildDistanceField"); return RBV_ERR_STATE; }
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_VEL, vel3, 3);
  TRY(stage_commit(c, slot));
  TRY(enqueue_minimize(c, old_map));
  return fetch_lm(c, vel3, Rvel9, score);
}

extern "C" int rbv_forward_match(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, int* n_matches) {
  CHECK_CTX(c);
  MAP_READY(c, old_map);
  MAP_READY(c, new_map);
  MapUse map_use_(c, old_map, new_map);
  if (!old_map || !new_map) return RBV_ERR_ARG;
  TRY(tr_forward_match(c, old_map, new_map));
  if (n_matches) {
    TRY(fetch_cnt(c));
    for (int b = 0; b < c->P.lanes; ++b) n_matches[b] = c->h_cnt[b * CNT_STRIDE + CNT_FWD];
  }
  return RBV_OK;
}

static void copy_extrot_out(rbv_ctx* c, const float* h_out, float* Wx36, float* X6, int* ok) {
  for (int b = 0; b < c->P.lanes; ++b) {
    const float* o = h_out + (size_t)b * OUT_STRIDE;
    if (Wx36) memcpy(Wx36 + b * 36, o + OUT_WX, 144);
    if (X6) memcpy(X6 + b * 6, o + OUT_X, 24);
    if (ok) ok[b] = o[OUT_OK] != 0.0f;
  }
}

extern "C" int rbv_ext_rot_vel(rbv_ctx* c, rbv_map* new_map, const float* vel3, float* Wx36, float* X6, int* ok) {
  CHECK_CTX(c);
  MAP_READY(c, new_map);
  MapUse map_use_(c, new_map);
  if (!new_map || !vel3) return RBV_ERR_ARG;
  if (c->field_map != new_map) { rbv_set_error("extRotVel: map is not the distance-field map (core.cpp:195)"); return RBV_ERR_STATE; }
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_VEL, vel3, 3);
  TRY(stage_commit(c, slot));
  TRY(tr_ext_rot_vel(c, new_map, 0));
  TRY(fetch_out(c));
  copy_extrot_out(c, c->h_out, Wx36, X6, ok);
  return RBV_OK;
}

static void put_directed(rbv_ctx* c, float* h, const float* vel3, const float* Rvel9, const float* Rback9, float max_radius) {
  for (int b = 0; b < c->P.lanes; ++b) {
    put(h, b, IN_VEL, vel3, 3); put(h, b, IN_RVEL, Rvel9, 9); put(h, b, IN_RBACK, Rback9, 9);
    h[(size_t)b * IN_STRIDE + IN_MAXRADIUS] = max_radius;
  }
}

extern "C" int rbv_directed_match(rbv_ctx* c, rbv_map* new_map, rbv_map* old_map, const float* vel3, const float* Rvel9,
                                  const float* Rback9, float max_radius, int* kf_matches, int* n_matches) {
  CHECK_CTX(c);
  MAP_READY(c, new_map);
  MAP_READY(c, old_map);
  MapUse map_use_(c, new_map, old_map);
  if (!new_map || !old_map || !vel3 || !Rvel9 || !Rback9) return RBV_ERR_ARG;
  int slot;
  float* h = stage_begin(c, &slot);
  put_directed(c, h, vel3, Rvel9, Rback9, max_radius);
  TRY(stage_commit(c, slot));
  TRY(tr_directed_match(c, new_map, old_map));
  if (kf_matches || n_matches) {
    TRY(fetch_cnt(c));
    for (int b = 0; b < c->P.lanes; ++b) {
      if (n_matches) n_matches[b] = c->h_cnt[b * CNT_STRIDE + CNT_DIR];
      if (kf_matches) kf_matches[b] = c->h_cnt[b * CNT_STRIDE + CNT_KF];
    }
  }
  return RBV_OK;
}

extern "C" int rbv_regularize(rbv_ctx* c, rbv_map* map, int* n_regularized) {
  CHECK_CTX(c);
  MAP_READY(c, map);
  MapUse map_use_(c, map);
  if (!map) return RBV_ERR_ARG;
  TRY(tr_regularize(c, map));
  if (n_regularized) {
    TRY(fetch_cnt(c));
    for (int b = 0; b < c->P.lanes; ++b) n_regularized[b] = c->h_cnt[b * CNT_STRIDE + CNT_REG];
  }
  return RBV_OK;
}

extern "C" int rbv_update_inverse_depth(rbv_ctx* c, const float* vel3) {
  CHECK_CTX(c);
  MAP_READY(c, c->field_map);
  MapUse map_use_(c, c->field_map);
  if (!vel3) return RBV_ERR_ARG;
  if (!c->field_map) { rbv_set_error("updateInverseDepth before buildDistanceField"); return RBV_ERR_STATE; }
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_VEL, vel3, 3);
  TRY(stage_commit(c, slot));
  return tr_kalman(c, c->field_map);
}

// ---- fused per-frame sequences (rebvio.cpp:142-177 and :223-259) ----
// first half of a tracker frame as one (graph-replayed) kernel sequence; R.T() in lane_in[IN_R], lane_in[IN_VEL] = 0
static int enqueue_g1(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map) {
  MAP_READY(c, old_map);
  MAP_READY(c, new_map);
  MapUse map_use_(c, old_map, new_map);
  const bool pre = new_map->field_built != 0;   // the field was built behind the map's detect: the kernels below read the map's own field
  FieldScope fs(c, new_map, pre);
  if (pre) c->field_map = new_map;              // what tr_build_field would have left (the stages below check it)
  return run_graphed(c, GK_TRACK1, old_map, new_map, pre ? 1 : 0, [&]() -> int {
    SeqFused fused(c, 0);
    // two independent branches: the distance field of the NEW map (the longest kernel of the half, instruction bound at ~45 %
    // occupancy) and rotation + sigma_rho quantile + LM seed of the OLD map; they meet in front of the first tryVel
    RBV_CUDA(cudaEventRecord(c->ev_fork, c->stream));
    RBV_CUDA(cudaStreamWaitEvent(c->side_stream, c->ev_fork, 0));
    {
      cudaStream_t main_stream = c->stream;
      c->stream = c->side_stream;
      int rc = tr_rotate(c, old_map);
      if (rc == RBV_OK) rc = enqueue_minimize(c, old_map, 1);
      c->stream = main_stream;
      TRY(rc);
    }
    RBV_CUDA(cudaEventRecord(c->ev_join, c->side_stream));
    if (!pre) TRY(tr_build_field(c, new_map));
    RBV_CUDA(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    TRY(enqueue_minimize(c, old_map, 2));
    TRY(tr_forward_match(c, old_map, new_map));
    return tr_ext_rot_vel(c, new_map, 1);
  });
}

// second half; R0, V, P_V, Rgva, max_radius in lane_in
static int enqueue_g2(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, int min_matches) {
  MAP_READY(c, old_map);
  MAP_READY(c, new_map);
  MapUse map_use_(c, old_map, new_map);
  return run_graphed(c, GK_TRACK2, old_map, new_map, min_matches, [&]() -> int {
    SeqFused fused(c, min_matches, new_map);   // lanes below the match floor skip regularise + Kalman (rebvio.cpp:247-252)
    TRY(tr_rotate(c, old_map));
    TRY(tr_directed_match(c, new_map, old_map));
    TRY(tr_regularize(c, new_map));
    return tr_kalman(c, new_map);
  });
}

extern "C" int rbv_track_g1(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, const float* R9, float* vel3, float* Rvel9, float* Wx36,
                            float* X6, int* ok) {
  CHECK_CTX(c);
  if (!old_map || !new_map || !R9) return RBV_ERR_ARG;
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_R, R9, 9);   // IN_VEL stays 0: imu_state_.Vg = Zeros (rebvio.cpp:168)
  TRY(stage_commit(c, slot));
  TRY(enqueue_g1(c, old_map, new_map));
  c->field_map = new_map;   // tr_build_field's host-side effect, also when the sequence was replayed
  RBV_CUDA(cudaMemcpyAsync(c->h_out, c->lane_out, sizeof(float) * (size_t)c->P.lanes * OUT_STRIDE, cudaMemcpyDeviceToHost, c->stream));
  TRY(fetch_lm(c, vel3, Rvel9, nullptr));
  copy_extrot_out(c, c->h_out, Wx36, X6, ok);
  return RBV_OK;
}

extern "C" int rbv_track_g2(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, const float* R0_9, const float* V3, const float* PV9,
                            const float* Rgva9, float max_radius, int min_matches, int* n_matches, int* kf_matches) {
  CHECK_CTX(c);
  if (!old_map || !new_map || !R0_9 || !V3 || !PV9 || !Rgva9) return RBV_ERR_ARG;
  if (c->field_map != new_map) { rbv_set_error("track_g2: new_map is not the distance-field map"); return RBV_ERR_STATE; }
  int slot;
  float* h = stage_begin(c, &slot);
  for (int b = 0; b < c->P.lanes; ++b) put(h, b, IN_R, R0_9, 9);
  put_directed(c, h, V3, PV9, Rgva9, max_radius);
  TRY(stage_commit(c, slot));
  TRY(enqueue_g2(c, old_map, new_map, min_matches));
  TRY(fetch_cnt(c));
  for (int b = 0; b < c->P.lanes; ++b) {
    if (n_matches) n_matches[b] = c->h_cnt[b * CNT_STRIDE + CNT_DIR];
    if (kf_matches) kf_matches[b] = c->h_cnt[b * CNT_STRIDE + CNT_KF];
  }
  return RBV_OK;
}

// ---- host block between the halves + whole-frame call ----
static void so3_exp_d(const double w[3], double R[9]) {   // Rodrigues; TooN::SO3::exp equivalent to double rounding
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
  double a, b;
  if (th < 1e-12) { a = 1.0; b = 0.0; }
  else { a = sin(th) / th; b = (1.0 - cos(th)) / th2; }
  const double K[9] = {0.0, -w[2], w[1], w[2], 0.0, -w[0], -w[1], w[0], 0.0};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double kk = 0.0;
      for (int k = 0; k < 3; ++k) kk += K[i * 3 + k] * K[k * 3 + j];
      R[i * 3 + j] = (i == j ? 1.0 : 0.0) + a * K[i * 3 + j] + b * kk;
    }
}
static bool inv6_d(const double* A, double* inv) {   // Gauss-Jordan with partial pivoting
  double M[6][12];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { M[i][j] = A[i * 6 + j]; M[i][6 + j] = i == j ? 1.0 : 0.0; }
  for (int c2 = 0; c2 < 6; ++c2) {
    int piv = c2;
    for (int r = c2 + 1; r < 6; ++r) if (fabs(M[r][c2]) > fabs(M[piv][c2])) piv = r;
    if (!(fabs(M[piv][c2]) > 0.0) || !isfinite(M[piv][c2])) return false;
    if (piv != c2) for (int j = 0; j < 12; ++j) { const double t = M[c2][j]; M[c2][j] = M[piv][j]; M[piv][j] = t; }
    const double d = 1.0 / M[c2][c2];
    for (int j = 0; j < 12; ++j) M[c2][j] *= d;
    for (int r = 0; r < 6; ++r) {
      if (r == c2) continue;
      const double f = M[r][c2];
      if (f != 0.0) for (int j = 0; j < 12; ++j) M[r][j] -= f * M[c2][j];
    }
  }
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) inv[i * 6 + j] = M[i][6 + j];
  return true;
}
extern "C" int rbv_visual_fusion(int lanes, const float* R9, const float* Vg3, const float* Wx36, const float* X6, float* R0_9, float* V3,
                                 float* PV9, float* Rgva9) {
  if (lanes < 1 || !R9 || !Vg3 || !Wx36 || !X6 || !R0_9 || !V3 || !PV9 || !Rgva9) { rbv_set_error("rbv_visual_fusion: null argument"); return RBV_ERR_ARG; }
  for (int b = 0; b < lanes; ++b) {
    const float* X = X6 + b * 6;
    const double w[3] = {X[3], X[4], X[5]};
    double R0[9], A[36], Ai[36];
    so3_exp_d(w, R0);
    for (int i = 0; i < 3; ++i) {
      double v = 0.0;
      for (int k = 0; k < 3; ++k) v += R0[i * 3 + k] * (double)Vg3[b * 3 + k];
      V3[b * 3 + i] = (float)(v + (double)X[i]);
    }
    for (int i = 0; i < 36; ++i) A[i] = Wx36[b * 36 + i];
    const bool ok = inv6_d(A, Ai);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) PV9[b * 9 + i * 3 + j] = ok ? (float)Ai[i * 6 + j] : (i == j ? 1e10f : 0.0f);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double t = 0.0;   // (R0 R^T)[j][i] -> Rgva[i][j]
        for (int k = 0; k < 3; ++k) t += R0[j * 3 + k] * (double)R9[b * 9 + i * 3 + k];
        Rgva9[b * 9 + i * 3 + j] = (float)t;
        R0_9[b * 9 + i * 3 + j] = (float)R0[i * 3 + j];
      }
  }
  return RBV_OK;
}

extern "C" int rbv_track(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, const float* R9, float max_radius, int min_matches, float* Vg3,
                         float* PVg9, float* Wx36, float* X6, int* ok, float* V3, int* n_matches, int* kf_matches, int* n_keylines_new,
                         float* auto_threshold_new) {
  CHECK_CTX(c);
  if (!old_map || !new_map || !R9) return RBV_ERR_ARG;
  const int B = c->P.lanes;
  std::vector<float> buf((size_t)B * (9 + 3 + 9 + 36 + 6 + 9 + 3 + 9 + 9));
  float* Rt = buf.data();
  float* vg = Rt + (size_t)B * 9;
  float* pvg = vg + (size_t)B * 3;
  float* wx = pvg + (size_t)B * 9;
  float* x = wx + (size_t)B * 36;
  float* r0 = x + (size_t)B * 6;
  float* v = r0 + (size_t)B * 9;
  float* pv = v + (size_t)B * 3;
  float* rg = pv + (size_t)B * 9;
  std::vector<int> okv((size_t)B);
  for (int b = 0; b < B; ++b)
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Rt[b * 9 + i * 3 + j] = R9[b * 9 + j * 3 + i];           // old->rotateKeylines(R.T()), rebvio.cpp:165
  TRY(rbv_track_g1(c, old_map, new_map, Rt, vg, pvg, wx, x, okv.data()));
  TRY(rbv_visual_fusion(B, R9, vg, wx, x, r0, v, pv, rg));
  TRY(rbv_track_g2(c, old_map, new_map, r0, v, pv, rg, max_radius, min_matches, n_matches, kf_matches));
  if (Vg3) memcpy(Vg3, vg, sizeof(float) * B * 3);
  if (PVg9) memcpy(PVg9, pvg, sizeof(float) * B * 9);
  if (Wx36) memcpy(Wx36, wx, sizeof(float) * B * 36);
  if (X6) memcpy(X6, x, sizeof(float) * B * 6);
  if (ok) memcpy(ok, okv.data(), sizeof(int) * B);
  if (V3) memcpy(V3, v, sizeof(float) * B * 3);
  for (int b = 0; b < B; ++b) {   // rode home with the match counters (k_rotate of the second half)
    if (n_keylines_new) n_keylines_new[b] = c->h_cnt[b * CNT_STRIDE + CNT_NKEY];
    if (auto_threshold_new) memcpy(auto_threshold_new + b, c->h_cnt + b * CNT_STRIDE + CNT_ATHR, sizeof(float));
  }
  return RBV_OK;
}

// ---- IMU fusion block (rbv_fusion.cu) ----
extern "C" void rbv_default_imu_cfg(rbv_imu_cfg* cfg) {   // types::ImuStateConfig, imu.hpp:153-168
  if (!cfg) return;
  *cfg = rbv_imu_cfg{9.81f, 2e-3f, 0.2e3f, 2.0e-3f, 1.6968e-04f, 1.9393e-05f, 1e-7f, 1e-2f, 1e-4f, 1.2e-3f, 1, 10, {0.0188f, 0.0037f, 0.0776f}};
}

extern "C" int rbv_imu_reset(rbv_ctx* c, const rbv_imu_cfg* cfg) {
  CHECK_CTX(c);
  if (!cfg) return RBV_ERR_ARG;
  return fu_reset(c, cfg);
}

static int imu_ready(rbv_ctx* c) {
  if (!c->imu_state) { rbv_set_error("IMU fusion call before rbv_imu_reset"); return RBV_ERR_STATE; }
  return RBV_OK;
}
static rbv_imu_in* imu_host_in(rbv_ctx* c) { return static_cast<rbv_imu_in*>(c->imu_host); }
static rbv_imu_out* imu_host_out(rbv_ctx* c) { return reinterpret_cast<rbv_imu_out*>(imu_host_in(c) + c->P.lanes); }
static float* imu_host_stage(rbv_ctx* c) { return reinterpret_cast<float*>(imu_host_out(c) + c->P.lanes); }

extern "C" int rbv_imu_prerotate(rbv_ctx* c, const rbv_imu_in* in, float* Rt9) {
  CHECK_CTX(c);
  if (!in) return RBV_ERR_ARG;
  TRY(imu_ready(c));
  const size_t B = c->P.lanes;
  RBV_CUDA(cudaStreamSynchronize(c->stream));   // the pinned mirror is reused every call
  memcpy(imu_host_in(c), in, sizeof(rbv_imu_in) * B);
  RBV_CUDA(cudaMemcpyAsync(c->imu_in, imu_host_in(c), sizeof(rbv_imu_in) * B, cudaMemcpyHostToDevice, c->stream));
  TRY(fu_pre(c));
  if (Rt9) {
    RBV_CUDA(cudaMemcpyAsync(imu_host_out(c), c->imu_out, sizeof(rbv_imu_out) * B, cudaMemcpyDeviceToHost, c->stream));
    RBV_CUDA(cudaStreamSynchronize(c->stream));
    for (size_t b = 0; b < B; ++b) memcpy(Rt9 + b * 9, imu_host_out(c)[b].Rt, 36);
  }
  return RBV_OK;
}

extern "C" int rbv_imu_fuse(rbv_ctx* c, const float* Vg3, const float* Wx36, const float* X6, rbv_imu_out* out) {
  CHECK_CTX(c);
  if (!Vg3 || !Wx36 || !X6 || !out) return RBV_ERR_ARG;
  TRY(imu_ready(c));
  const size_t B = c->P.lanes;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  float* h = imu_host_stage(c);
  for (size_t b = 0; b < B; ++b) {
    memcpy(h + b * 48, Vg3 + b * 3, 12);
    memcpy(h + b * 48 + 3, Wx36 + b * 36, 144);
    memcpy(h + b * 48 + 39, X6 + b * 6, 24);
  }
  RBV_CUDA(cudaMemcpyAsync(c->imu_stage, h, sizeof(float) * 48 * B, cudaMemcpyHostToDevice, c->stream));
  float max_radius = 0.0f;   // the stage call leaves lane_in[IN_MAXRADIUS] to rbv_directed_match / rbv_track_g2
  TRY(fu_fuse(c, c->imu_stage, 48, c->imu_stage + 3, 48, c->imu_stage + 39, 48, max_radius));
  RBV_CUDA(cudaMemcpyAsync(imu_host_out(c), c->imu_out, sizeof(rbv_imu_out) * B, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  memcpy(out, imu_host_out(c), sizeof(rbv_imu_out) * B);
  return RBV_OK;
}

extern "C" int rbv_track_imu(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map, const rbv_imu_in* in, float max_radius,
                             int min_matches, float* Vg3, float* PVg9, float* Wx36, float* X6, int* ok, rbv_imu_out* out,
                             int* n_matches, int* kf_matches, int* n_keylines_new, float* auto_threshold_new) {
  if (!c) { rbv_set_error("null context"); return RBV_ERR_ARG; }
  // outermost entry (never called with the lock held): the lock is dropped while the call waits for the device at its end, so that the
  // reference's acquisition thread can enqueue the next detect on the front-end stream meanwhile (RBV_TUNE_FE_STREAM)
  std::unique_lock<std::recursive_mutex> ctx_lock_(*static_cast<std::recursive_mutex*>(c->mtx));
  RBV_CUDA(cudaSetDevice(c->device));
  if (!old_map || !new_map || !in) return RBV_ERR_ARG;
  TRY(imu_ready(c));
  const size_t B = c->P.lanes;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  memcpy(imu_host_in(c), in, sizeof(rbv_imu_in) * B);
  RBV_CUDA(cudaMemcpyAsync(c->imu_in, imu_host_in(c), sizeof(rbv_imu_in) * B, cudaMemcpyHostToDevice, c->stream));
  TRY(fu_pre(c));
  TRY(enqueue_g1(c, old_map, new_map));
  c->field_map = new_map;
  TRY(fu_fuse(c, c->lm[0].vel, (int)(sizeof(LMState) / sizeof(float)), c->lane_out + OUT_WX, OUT_STRIDE, c->lane_out + OUT_X,
              OUT_STRIDE, max_radius));
  // the tracker's first-half results leave the device before the second half may overwrite them (same stream: ordered)
  RBV_CUDA(cudaMemcpyAsync(trk_out(c), c->lane_out, sizeof(float) * B * OUT_STRIDE, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaMemcpyAsync(trk_lm(c), c->lm, sizeof(LMState) * B, cudaMemcpyDeviceToHost, c->stream));
  TRY(enqueue_g2(c, old_map, new_map, min_matches));
  RBV_CUDA(cudaMemcpyAsync(trk_imu(c), c->imu_out, sizeof(rbv_imu_out) * B, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaMemcpyAsync(trk_cnt(c), c->lane_cnt, sizeof(int) * B * CNT_STRIDE, cudaMemcpyDeviceToHost, c->stream));
  if (c->tune[RBV_TUNE_FE_STREAM] == 1) {
    RBV_CUDA(cudaEventRecord(old_map->used, c->stream)); old_map->used_valid = 1;
    RBV_CUDA(cudaEventRecord(new_map->used, c->stream)); new_map->used_valid = 1;
  }
  {
    cudaStream_t st = c->stream;
    // the results land in h_trk, which no other entry point touches (one estimation thread per context, like the reference's): the acquisition
    // thread's detect / map_size / threshold calls use h_misc and the map's own h_fe while the lock is dropped
    const bool drop = c->tune[RBV_TUNE_FE_STREAM] == 1;
    if (drop) ctx_lock_.unlock();
    const cudaError_t se = cudaStreamSynchronize(st);
    if (drop) { ctx_lock_.lock(); cudaSetDevice(c->device); }
    RBV_CUDA(se);
  }
  const LMState* L = trk_lm(c);
  const int* hc = trk_cnt(c);
  for (size_t b = 0; b < B; ++b) {
    if (Vg3) memcpy(Vg3 + b * 3, L[b].vel, 12);
    if (PVg9) memcpy(PVg9 + b * 9, L[b].Rvel, 36);
    if (n_matches) n_matches[b] = hc[b * CNT_STRIDE + CNT_DIR];
    if (kf_matches) kf_matches[b] = hc[b * CNT_STRIDE + CNT_KF];
    if (n_keylines_new) n_keylines_new[b] = hc[b * CNT_STRIDE + CNT_NKEY];
    if (auto_threshold_new) memcpy(auto_threshold_new + b, hc + b * CNT_STRIDE + CNT_ATHR, sizeof(float));
  }
  copy_extrot_out(c, trk_out(c), Wx36, X6, ok);
  if (out) memcpy(out, trk_imu(c), sizeof(rbv_imu_out) * B);
  return RBV_OK;
}

// ---- host mirror (EdgeMap::keylines(), mask(), size(), threshold(); edge_map.hpp:40-75) ----
extern "C" int rbv_map_size(rbv_ctx* c, rbv_map* m, int* n) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || !n) return RBV_ERR_ARG;
  RBV_CUDA(cudaMemcpyAsync(c->h_misc, m->count, sizeof(int) * c->P.lanes, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  memcpy(n, c->h_misc, sizeof(int) * c->P.lanes);
  return RBV_OK;
}
extern "C" int rbv_map_threshold(rbv_ctx* c, rbv_map* m, float* t) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || !t) return RBV_ERR_ARG;
  RBV_CUDA(cudaMemcpyAsync(c->h_misc, m->threshold, sizeof(float) * c->P.lanes, cudaMemcpyDeviceToHost, c->stream));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  memcpy(t, c->h_misc, sizeof(float) * c->P.lanes);
  return RBV_OK;
}
extern "C" int rbv_map_set_threshold(rbv_ctx* c, rbv_map* m, const float* t) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || !t) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, m->threshold, t, sizeof(float) * c->P.lanes, cudaMemcpyHostToDevice));
  return RBV_OK;
}

template <typename T>
static cudaError_t pull(rbv_ctx* c, std::vector<T>& h, const T* d, size_t lane, size_t K, int n) {
  h.resize(n > 0 ? n : 1);
  return copy_sync(c, h.data(), d + lane * K, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost);
}
template <typename T>
static cudaError_t push(rbv_ctx* c, const std::vector<T>& h, T* d, size_t lane, size_t K, int n) {
  return copy_sync(c, d + lane * K, h.data(), sizeof(T) * (size_t)n, cudaMemcpyHostToDevice);
}

extern "C" int rbv_map_download(rbv_ctx* c, rbv_map* m, int lane, rbv_keyline* kl, uint32_t* pixel_idx, int cap, int* n_out) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || lane < 0 || lane >= c->P.lanes || !n_out) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  int n = 0;
  RBV_CUDA(copy_sync(c, &n, m->count + lane, sizeof(int), cudaMemcpyDeviceToHost));
  *n_out = n;
  if (n > cap) { rbv_set_error("rbv_map_download: capacity %d < %d keylines", cap, n); return RBV_ERR_ARG; }
  if (n == 0) return RBV_OK;
  const size_t K = c->P.kmax, L = lane;
  std::vector<float2> pos, pi, mpi, g, mg;
  std::vector<float> gn, mgn, rho, sr;
  std::vector<int> ip, in, mi, mf, mk;
  std::vector<unsigned> mt, px;
  RBV_CUDA(pull(c, pos, m->k.pos, L, K, n)); RBV_CUDA(pull(c, pi, m->k.pos_img, L, K, n)); RBV_CUDA(pull(c, mpi, m->k.match_pos_img, L, K, n));
  RBV_CUDA(pull(c, g, m->k.grad, L, K, n)); RBV_CUDA(pull(c, mg, m->k.match_grad, L, K, n));
  RBV_CUDA(pull(c, gn, m->k.gnorm, L, K, n)); RBV_CUDA(pull(c, mgn, m->k.match_gnorm, L, K, n)); RBV_CUDA(pull(c, rho, m->k.rho, L, K, n));
  RBV_CUDA(pull(c, sr, m->k.srho, L, K, n));
  RBV_CUDA(pull(c, ip, m->k.id_prev, L, K, n)); RBV_CUDA(pull(c, in, m->k.id_next, L, K, n)); RBV_CUDA(pull(c, mi, m->k.match_id, L, K, n));
  RBV_CUDA(pull(c, mf, m->k.match_fwd, L, K, n)); RBV_CUDA(pull(c, mk, m->k.match_kf, L, K, n));
  RBV_CUDA(pull(c, mt, m->k.matches, L, K, n)); RBV_CUDA(pull(c, px, m->k.pixel, L, K, n));
  for (int i = 0; i < n; ++i) {
    if (kl) {
      rbv_keyline& k = kl[i];
      k.pos[0] = pos[i].x; k.pos[1] = pos[i].y; k.pos_img[0] = pi[i].x; k.pos_img[1] = pi[i].y;
      k.match_pos_img[0] = mpi[i].x; k.match_pos_img[1] = mpi[i].y; k.gradient[0] = g[i].x; k.gradient[1] = g[i].y;
      k.match_gradient[0] = mg[i].x; k.match_gradient[1] = mg[i].y; k.gradient_norm = gn[i]; k.match_gradient_norm = mgn[i];
      k.rho = rho[i]; k.sigma_rho = sr[i];
      k.id = -1;  // KeyLine::id is never assigned by the reference (keyline.hpp:52)
      k.id_prev = ip[i]; k.id_next = in[i]; k.match_id = mi[i]; k.match_id_forward = mf[i]; k.match_id_keyframe = mk[i];
      k.matches = mt[i];
    }
    if (pixel_idx) pixel_idx[i] = px[i];
  }
  return RBV_OK;
}

extern "C" int rbv_map_upload(rbv_ctx* c, rbv_map* m, int lane, const rbv_keyline* kl, const uint32_t* pixel_idx, int n) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || lane < 0 || lane >= c->P.lanes || n < 0 || n > c->P.keylines_max || (n && !kl)) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  const size_t K = c->P.kmax, L = lane, N = c->P.npix;
  std::vector<float2> pos(n), pi(n), mpi(n), g(n), mg(n);
  std::vector<float> gn(n), mgn(n), rho(n), sr(n);
  std::vector<int> ip(n), in(n), mi(n), mf(n), mk(n);
  std::vector<unsigned> mt(n);
  for (int i = 0; i < n; ++i) {
    const rbv_keyline& k = kl[i];
    pos[i] = make_float2(k.pos[0], k.pos[1]); pi[i] = make_float2(k.pos_img[0], k.pos_img[1]);
    mpi[i] = make_float2(k.match_pos_img[0], k.match_pos_img[1]); g[i] = make_float2(k.gradient[0], k.gradient[1]);
    mg[i] = make_float2(k.match_gradient[0], k.match_gradient[1]); gn[i] = k.gradient_norm; mgn[i] = k.match_gradient_norm;
    rho[i] = k.rho; sr[i] = k.sigma_rho; ip[i] = k.id_prev; in[i] = k.id_next; mi[i] = k.match_id; mf[i] = k.match_id_forward;
    mk[i] = k.match_id_keyframe; mt[i] = k.matches;
  }
  if (n) {
    RBV_CUDA(push(c, pos, m->k.pos, L, K, n)); RBV_CUDA(push(c, pi, m->k.pos_img, L, K, n)); RBV_CUDA(push(c, mpi, m->k.match_pos_img, L, K, n));
    RBV_CUDA(push(c, g, m->k.grad, L, K, n)); RBV_CUDA(push(c, mg, m->k.match_grad, L, K, n));
    RBV_CUDA(push(c, gn, m->k.gnorm, L, K, n)); RBV_CUDA(push(c, mgn, m->k.match_gnorm, L, K, n)); RBV_CUDA(push(c, rho, m->k.rho, L, K, n));
    RBV_CUDA(push(c, sr, m->k.srho, L, K, n));
    RBV_CUDA(push(c, ip, m->k.id_prev, L, K, n)); RBV_CUDA(push(c, in, m->k.id_next, L, K, n)); RBV_CUDA(push(c, mi, m->k.match_id, L, K, n));
    RBV_CUDA(push(c, mf, m->k.match_fwd, L, K, n)); RBV_CUDA(push(c, mk, m->k.match_kf, L, K, n)); RBV_CUDA(push(c, mt, m->k.matches, L, K, n));
  }
  RBV_CUDA(copy_sync(c, m->count + lane, &n, sizeof(int), cudaMemcpyHostToDevice));
  if (pixel_idx) {
    std::vector<int> mask(N, -1);
    std::vector<unsigned> px(n > 0 ? n : 1);
    for (int i = 0; i < n; ++i) {
      if (pixel_idx[i] >= N) { rbv_set_error("rbv_map_upload: pixel index out of range"); return RBV_ERR_ARG; }
      px[i] = pixel_idx[i];
      if (mask[pixel_idx[i]] < 0) mask[pixel_idx[i]] = i;   // unordered_map::emplace keeps the first
    }
    RBV_CUDA(copy_sync(c, m->mask + L * N, mask.data(), sizeof(int) * N, cudaMemcpyHostToDevice));
    if (n) RBV_CUDA(push(c, px, m->k.pixel, L, K, n));
  }
  if (c->field_map == m) { FieldScope fs(c, m, m->field_built != 0); TRY(tr_pack_targets(c, m)); }   // tryVel reads a packed copy of the field map's keylines
  return RBV_OK;
}

// ---- debug taps ----
extern "C" int rbv_debug_read_image(rbv_ctx* c, int lane, int which, float* out) {
  CHECK_CTX(c);
  if (lane < 0 || lane >= c->P.lanes || which < 0 || which > 4 || !out) return RBV_ERR_ARG;
  const size_t N = c->P.npix;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  if (which == 4) {   // the float frame the front end worked on (after the u8 conversion / undistortion of the host-frame paths)
    RBV_CUDA(copy_sync(c, out, c->img + (size_t)lane * N, sizeof(float) * N, cudaMemcpyDeviceToHost));
    return RBV_OK;
  }
  if (which < 2) {
    RBV_CUDA(copy_sync(c, out, c->bufA + ((size_t)lane * 2 + which) * N, sizeof(float) * N, cudaMemcpyDeviceToHost));
    return RBV_OK;
  }
  float* tmp = c->rowsum;  // scratch
  TRY(fe_dog_mag_debug(c, lane, which, tmp));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, out, tmp, sizeof(float) * N, cudaMemcpyDeviceToHost));
  return RBV_OK;
}

extern "C" int rbv_debug_read_mask(rbv_ctx* c, rbv_map* m, int lane, int* out) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || lane < 0 || lane >= c->P.lanes || !out) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, out, m->mask + (size_t)lane * c->P.npix, sizeof(int) * (size_t)c->P.npix, cudaMemcpyDeviceToHost));
  return RBV_OK;
}

extern "C" int rbv_debug_read_field(rbv_ctx* c, int lane, int* id, int* distance) {
  CHECK_CTX(c);
  if (lane < 0 || lane >= c->P.lanes) return RBV_ERR_ARG;
  const size_t N = c->P.npix;
  std::vector<unsigned> f(N);
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, f.data(), c->field + (size_t)lane * N, sizeof(unsigned) * N, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; ++i) {
    const bool empty = f[i] == RBV_FIELD_EMPTY;
    if (id) id[i] = empty ? -1 : (int)(0xFFFFFFu - (f[i] & 0xFFFFFFu));
    if (distance) distance[i] = empty ? -1 : (int)(f[i] >> 24);
  }
  return RBV_OK;
}

extern "C" int rbv_debug_read_residuals(rbv_ctx* c, rbv_map* m, int lane, float* out, int cap) {
  CHECK_CTX(c);
  MAP_READY(c, m);
  MapUse map_use_(c, m);
  if (!m || lane < 0 || lane >= c->P.lanes || !out) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  int n = 0;
  RBV_CUDA(copy_sync(c, &n, m->count + lane, sizeof(int), cudaMemcpyDeviceToHost));
  if (n > cap) n = cap;
  if (n) RBV_CUDA(copy_sync(c, out, m->k.residual + (size_t)lane * c->P.kmax, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
  return RBV_OK;
}

extern "C" int rbv_debug_box_pass(rbv_ctx* c, int scale, int pass, const float* in, float* out) {
  CHECK_CTX(c);
  if (scale < 0 || scale > 1 || pass < 0 || pass > 2 || !in || !out) return RBV_ERR_ARG;
  const size_t N = c->P.npix;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, c->img, in, sizeof(float) * N, cudaMemcpyHostToDevice));
  TRY(fe_box_pass_debug(c, scale, pass, c->img, c->bufB));
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, out, c->bufB, sizeof(float) * N, cudaMemcpyDeviceToHost));
  return RBV_OK;
}

extern "C" int rbv_debug_read_setup(rbv_ctx* c, float* pinv75, int* widths6) {
  if (!c) return RBV_ERR_ARG;
  if (pinv75) memcpy(pinv75, c->P.pinv, sizeof(float) * 75);
  if (widths6) for (int s = 0; s < 2; ++s) for (int p = 0; p < 3; ++p) widths6[s * 3 + p] = c->widths[s][p];
  return RBV_OK;
}

extern "C" int rbv_debug_read_lm_trace(rbv_ctx* c, int lane, float* out, int cap_floats) {
  CHECK_CTX(c);
  if (lane < 0 || lane >= c->P.lanes || !out) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  LMState L;
  RBV_CUDA(copy_sync(c, &L, c->lm + lane, sizeof(LMState), cudaMemcpyDeviceToHost));
  int n = (c->P.iterations + 1) * 16;
  if (n > cap_floats) n = cap_floats;
  memcpy(out, L.trace, sizeof(float) * n);
  return RBV_OK;
}

extern "C" int rbv_debug_read_probe_count(rbv_ctx* c, int lane, unsigned long long* probes) {
  CHECK_CTX(c);
  if (lane < 0 || lane >= c->P.lanes || !probes) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  RBV_CUDA(copy_sync(c, probes, c->probes + lane, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return RBV_OK;
}

// debug tap: one row of the ORDERED-mode addends (tryVel: planes 0..9, extRotVel: 100..126) of one lane, as left by the last evaluation kernel;
// plane -1: the chain kernel's per-chunk running sums of accumulator 0 (written when RBV_DEBUG_CHAIN is set in the environment)
extern "C" int rbv_debug_read_addends(rbv_ctx* c, int plane, int lane, float* out, int n) {
  CHECK_CTX(c);
  const bool ok_plane = plane == -1 || (plane >= 0 && plane < 10) || (plane >= 100 && plane < 127);
  if (!ok_plane || lane < 0 || lane >= c->P.lanes || !out || n < 0 || n > c->P.kmax) return RBV_ERR_ARG;
  RBV_CUDA(cudaStreamSynchronize(c->stream));
  if (plane >= 0) return tr_debug_addends(c, plane, lane, out, n);
  if (n) RBV_CUDA(copy_sync(c, out, c->fiabs + (size_t)lane * c->P.kmax, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
  return RBV_OK;
}
