/*
 * rebvio_b200.h — C ABI of the B200-native (sm_100a) REBVIO hot path.
 *
 * Drop-in boundary for the reference's C++ class API (baumlin/rebvio). The reference has no FFI layer; the
 * seam is the set of member functions that Rebvio's two worker loops call (rebvio/src/rebvio.cpp:70,142,165,
 * 169,172,177,223,232,245,256,259). Each entry point below names the member function it replaces. The C++
 * adapter classes in rebvio_b200/cpp/rebvio_b200/ carry the reference signatures and forward here.
 *
 * Conventions
 *   - plain C: opaque handles, POD structs, caller-owned host buffers, int status returns (0 = RBV_OK).
 *   - no exceptions / STL / torch types cross this boundary. All device memory is owned by the context.
 *   - a context holds B independent "lanes" (one lane = one image sequence). Every call processes all B
 *     lanes in one set of kernel launches; per-lane scalars are passed as arrays of length B (lane-major).
 *     B = 1 is the reference's single-stream shape.
 *   - host pointers may be pageable or pinned; calls are asynchronous on the context's CUDA stream unless
 *     they return data to the host, in which case they synchronise that stream before returning.
 *   - thread safety: every call that takes an rbv_ctx* holds that context's lock while it runs, so the reference's
 *     acquisition and estimation threads (rebvio.cpp:98-99) may share one context; calls are serialised in arrival
 *     order. rbv_last_error() is per thread. rbv_destroy must not race with other calls on the same context.
 *   - there is NO CPU fallback: if no sm_100 device / kernel image is usable, rbv_create fails with
 *     RBV_ERR_CUDA and rbv_last_error() says why.
 */
#ifndef REBVIO_B200_H_
#define REBVIO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RBV_OK 0
#define RBV_ERR_ARG 1      /* bad argument (null handle, lane count mismatch, size out of range) */
#define RBV_ERR_CUDA 2     /* CUDA runtime / driver error, no device, wrong architecture          */
#define RBV_ERR_STATE 3    /* call order violated (e.g. tracker stage before a distance field)   */

/* reduction mode for the tracker's J^T J / J^T F sums (core.cpp:130-139,244-245)                     */
#define RBV_REDUCE_FAST 0     /* warp-shuffle tree, FP64 accumulate, fixed order (deterministic; NOT the reference's bits) */
#define RBV_REDUCE_ORDERED 1  /* DEFAULT: per-accumulator serial float32 adds in keyline order == CPU bits (one chain warp  */
                              /* per lane behind the evaluation kernel, rbv_tracker.cu)                                  */

/* ---- POD mirrors of the reference's structs ------------------------------------------------------- */

/* rebvio::Camera public fields read by the hot path (camera.hpp:61-72) */
typedef struct rbv_camera {
  float fm, cx, cy;
  int rows, cols;
} rbv_camera;

/* rebvio::EdgeDetectorConfig (edge_detector.hpp:19-32); plane_fit_size=2 and num_bins=100 are constexpr there */
typedef struct rbv_detector_cfg {
  int keylines_ref;        /* 12000 */
  int keylines_max;        /* 16000 */
  float pos_neg_threshold; /* 0.4 */
  float dog_threshold;     /* 0.095259868922420 */
  float threshold;         /* 0.01 */
  float gain;              /* 5e-7 */
  float max_threshold;     /* 0.5 */
  float min_threshold;     /* 0.005 */
} rbv_detector_cfg;

/* rebvio::EdgeMapConfig (edge_map.hpp:19-26) */
typedef struct rbv_map_cfg {
  float pixel_uncertainty_match;  /* 2.0 */
  float match_threshold_norm;     /* 1.0 */
  float match_threshold_angle;    /* 45.0 */
  float regularization_threshold; /* 0.5 */
} rbv_map_cfg;

/* rebvio::CoreConfig (core.hpp:82-95) */
typedef struct rbv_core_cfg {
  float search_range;                        /* 40 */
  float reweight_distance;                   /* 2 */
  float match_treshold;                      /* 0.5 (spelling as in the reference) */
  unsigned int min_match_threshold;          /* 0 */
  unsigned int iterations;                   /* 5 */
  unsigned int global_min_matches_threshold; /* 500 */
  float pixel_uncertainty;                   /* 1 */
  float quantile_cutoff;                     /* 0.9 */
  int quantile_num_bins;                     /* 100 */
  float reshape_q_abs;                       /* 1e-4 */
} rbv_core_cfg;

/* rebvio::types::KeyLine (types/keyline.hpp:24-60), 84 bytes, same field order */
typedef struct rbv_keyline {
  float pos[2];
  float pos_img[2];
  float match_pos_img[2];
  float gradient[2];
  float match_gradient[2];
  float gradient_norm;
  float match_gradient_norm;
  float rho;
  float sigma_rho;
  int id;
  int id_prev;
  int id_next;
  int match_id;
  int match_id_forward;
  int match_id_keyframe;
  unsigned int matches;
} rbv_keyline;

void rbv_default_camera(rbv_camera* cam);             /* camera.hpp:26-37 (EuRoC cam0, 752x480) */
void rbv_default_detector_cfg(rbv_detector_cfg* cfg); /* edge_detector.hpp:19-32 */
void rbv_default_map_cfg(rbv_map_cfg* cfg);           /* edge_map.hpp:19-26 */
void rbv_default_core_cfg(rbv_core_cfg* cfg);         /* core.hpp:82-95 */

/* ---- handles ---------------------------------------------------------------------------------------- */

typedef struct rbv_ctx rbv_ctx; /* EdgeDetector + Core for B lanes: scale-space scratch, id mask, distance field */
typedef struct rbv_map rbv_map; /* B EdgeMaps (one per lane): keyline SoA + pixel->id mask, device resident    */

/* replaces EdgeDetector::EdgeDetector (edge_detector.cpp:17-26) + Core::Core (core.cpp:18-25) for `lanes`
 * sequences on CUDA device `device`. Pre-computes the box widths / divisor tables (scale_space.cpp:14-41,
 * 130-171) and the 3x25 plane-fit pseudo inverse (edge_detector.cpp:55-68). */
int rbv_create(const rbv_camera* cam, const rbv_detector_cfg* det, const rbv_map_cfg* map,
               const rbv_core_cfg* core, int device, int lanes, rbv_ctx** out);
void rbv_destroy(rbv_ctx* ctx);
const char* rbv_last_error(void);
int rbv_sync(rbv_ctx* ctx);
int rbv_set_reduce_mode(rbv_ctx* ctx, int mode); /* RBV_REDUCE_ORDERED (default, bit-exact) | RBV_REDUCE_FAST */
int rbv_lanes(const rbv_ctx* ctx);
/* run every subsequent call on the caller's CUDA stream (a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream)
 * instead of the context's own; lets a host framework time the work with its own events. */
int rbv_set_stream(rbv_ctx* ctx, void* cuda_stream);
/* number of kernels this context has launched since rbv_create (bench.py's gpu_launches) */
unsigned long long rbv_launch_count(const rbv_ctx* ctx);

/* per-kernel accounting. Every device operation of this library is one of rbv_kernel_count() named kernels.
 * rbv_profile_set: mode -1 = off (default), -2 = bracket every launch with CUDA events on the context's stream,
 * k >= 0 = bracket only launches of kernel k (cheap enough to leave on inside a timed region).
 * rbv_profile_read: summed event time of the bracketed launches of kernel `kid`, how many were bracketed, and how
 * many launches of it happened since rbv_profile_reset (synchronises the stream). */
int rbv_kernel_count(void);
const char* rbv_kernel_name(int kid);
int rbv_profile_set(rbv_ctx* ctx, int mode);
int rbv_profile_reset(rbv_ctx* ctx);
int rbv_profile_read(rbv_ctx* ctx, int kid, double* total_ms, unsigned long long* timed_launches,
                     unsigned long long* launches);

/* Camera::undistort on the device (camera.hpp:54-58 = cv::undistort(img, K_, D_), called from Rebvio::imageCallback,
 * rebvio.cpp:45, on the float32 x 3 frame). K4 = (fx, fy, cx, cy) - the reference passes (fm, fm, cx, cy), camera.hpp:39 -
 * and D5 = (k1, k2, p1, p2, k3), both float32 like the reference's K_ / D_ matrices. Once set, rbv_detect, rbv_detect_u8
 * and rbv_prefetch_frames take RAW camera frames and the library runs the fixed-point bilinear remap of cv::undistort
 * (INTER_BITS = 5, BORDER_CONSTANT 0) in front of the scale space; rbv_detect_device is unaffected (device frames are
 * taken as already undistorted). K4 = NULL switches it off. Bit-exact against cv2.undistort on the golden vectors.      */
int rbv_set_undistort(rbv_ctx* ctx, const float* K4, const float* D5);

/* implementation choices, for tests and profiling; 0 = automatic (the default) everywhere.
 *   RBV_TUNE_FIELD: distance field — 1 = one thread per keyline with global atomics, 2 = shared-memory row bands
 *   RBV_TUNE_BOX:   box passes of the scale space — 1 = row-prefix kernel + column/box kernel, 2 = fused kernel, first form (k_boxpass),
 *                   3 = the shared-memory-resident pair for few planes (k_rowscan_lat + k_colbox_lat), 4 = fused kernel, second form
 *                   (k_boxpass2: register delay lines, three TMA chunk buffers; box widths 7 / 9, cols <= 832, cols % 4 == 0)
 *   RBV_TUNE_TRYVEL: keylines per thread of the tryVel kernel — 1 (FAST mode only), 2 or 4
 *   RBV_TUNE_FIELD_ROWS: image rows per band of the shared-memory distance field (0 = automatic)
 *   RBV_TUNE_GRAPH: 1 = every kernel launched on the stream, 2 (default) = the kernel sequences of rbv_detect*, rbv_track_g1 and
 *                   rbv_track_g2 are captured once per (map pair) as CUDA graphs and replayed with one launch each
 *                   (one block per image plane, TMA-staged row chunks; needs cols % 4 == 0 and cols <= 832)
 *   RBV_TUNE_FE_STREAM: 1 = rbv_detect* enqueue on the context's own front-end stream (SURVEY.md 8(b) "Threading": the
 *                   reference detects frame n+1 on one thread while the other tracks frames n-1, n, rebvio.cpp:56-90 /
 *                   92-293). A detect without outputs returns at once, rbv_detect_wait collects its counts; tracker and
 *                   mirror calls wait on the device for the detects of the maps they are given. Results are identical.    */
#define RBV_TUNE_FIELD 0
#define RBV_TUNE_BOX 1
#define RBV_TUNE_TRYVEL 2
#define RBV_TUNE_GRAPH 3
#define RBV_TUNE_FIELD_ROWS 4
#define RBV_TUNE_FE_STREAM 5
#define RBV_TUNE_FIELD_IN_FE 7 /* with RBV_TUNE_FE_STREAM: Core::buildDistanceField of a map runs behind its detect on the front-end stream, into the map's own field */
/* key 6 is unused (was RBV_TUNE_BOX_SPLIT, measured and rejected in round 1) */
#define RBV_TUNE_ORDERED 8 /* RBV_REDUCE_ORDERED tryVel and extRotVel: 0 = one launch each: evaluation blocks + one chain CTA per lane adding behind them (addends staged through shared memory by bulk copies), 2 = the same chains as kernels of their own behind the evaluation, 3 = tryVel as 0, extRotVel as 2, 4 = as 0 with all iterations + 1 tryVel calls of Core::minimizeVel in ONE launch (k_minimize_ord*: the chain CTAs stay, every lane advances on its own; measured no faster: the slowest lane's chain is the floor either way), 1 = the first form (adding block per lane reading global memory) */
#define RBV_TUNE_COUNT 9
int rbv_set_tuning(rbv_ctx* ctx, int key, int value);

/* allocate / free a B-lane map (replaces std::make_shared<EdgeMap>, edge_detector.cpp:50). */
int rbv_map_create(rbv_ctx* ctx, rbv_map** out);
void rbv_map_release(rbv_ctx* ctx, rbv_map* map);

/* ---- front end -------------------------------------------------------------------------------------- */

/* replaces EdgeDetector::detect → buildEdgeMap + joinEdges + tuneThreshold (edge_detector.cpp:30-186) for B
 * frames. `images`: B pointers to host float32 frames (row pitch `pitch_bytes`), values 0..765 as produced by
 * Rebvio::imageCallback (rebvio.cpp:43-45). `threshold[B]` is the (already servoed) config_->threshold of each
 * lane (edge_detector.cpp:33-36 stays on the host: rbv_servo_threshold). Outputs (host, length B, optional):
 * keyline counts and the auto threshold written to map->threshold(). The call copies H2D, launches, and - when an output
 * pointer is given - returns after the counts are back on the host; with both outputs NULL (or on the front-end stream,
 * RBV_TUNE_FE_STREAM) it returns once the frames have been UPLOADED: the caller's buffers (pageable or pinned) may be reused as
 * soon as the call returns, the kernels may still be running (rbv_detect_wait / rbv_sync wait for those). */
int rbv_detect(rbv_ctx* ctx, rbv_map* map, const float* const* images, size_t pitch_bytes,
               const uint64_t* ts_us, const float* threshold, int* n_keylines, float* auto_threshold);

/* same for 8-bit frames as the camera driver delivers them: replaces Rebvio::imageCallback's conversion
 * `convertTo(CV_32FC1, 3.0)` (rebvio.cpp:43) + EdgeDetector::detect. The bytes are uploaded as they are (a quarter of
 * the float traffic) and multiplied by 3.0f on the device, which is exact. Undistortion (rebvio.cpp:45, cv::undistort)
 * is not part of this path: frames are taken as already rectified. */
int rbv_detect_u8(rbv_ctx* ctx, rbv_map* map, const uint8_t* const* images, size_t pitch_bytes,
                  const uint64_t* ts_us, const float* threshold, int* n_keylines, float* auto_threshold);

/* start uploading the NEXT frames (B host pointers, u8 if is_u8 else float32) on the context's copy stream and return
 * at once; the upload overlaps whatever runs on the compute stream (the tracker of the current frame — the reference
 * overlaps acquisition and estimation with two threads, rebvio.cpp:56-117). The next rbv_detect / rbv_detect_u8 call
 * whose images[0] equals this call's images[0] consumes the staged frames instead of copying again. Three slots. The
 * host buffers must stay valid and unchanged until that detect call returns (it waits for the staged upload to have finished
 * reading them before it returns). */
int rbv_prefetch_frames(rbv_ctx* ctx, const void* const* images, size_t pitch_bytes, int is_u8);

/* same, frames already resident in device memory (lane-major contiguous [B][rows][cols] float32). */
int rbv_detect_device(rbv_ctx* ctx, rbv_map* map, const float* d_images, const uint64_t* ts_us,
                      const float* threshold, int* n_keylines, float* auto_threshold);

/* Threshold servo on the device: every rbv_detect* call accepts threshold = NULL, meaning "servo config_->threshold of every lane
 * from the keyline counts of this context's previous detect" (edge_detector.cpp:33-36, the same float32 statement as
 * rbv_servo_threshold; the state starts at rbv_detector_cfg::threshold with a count of 0, and an explicit threshold array
 * overwrites it). Such a detect needs nothing read back from the previous one, so front ends can be enqueued frames ahead of
 * the tracker (RBV_TUNE_FE_STREAM); rbv_detect_wait's threshold_used returns what the servo chose.
 *
 * keyline counts, auto thresholds and thresholds used (all optional, length B) of the most recent detect into `map`: after a detect on the
 * front-end stream (RBV_TUNE_FE_STREAM) this waits for that detect only; otherwise it reads the map's counters. */
int rbv_detect_wait(rbv_ctx* ctx, rbv_map* map, int* n_keylines, float* auto_threshold, float* threshold_used);

/* host scalar servo of edge_detector.cpp:33-36: returns the new config_->threshold. */
float rbv_servo_threshold(const rbv_detector_cfg* cfg, float threshold, int last_keyline_count);

/* ---- tracker ---------------------------------------------------------------------------------------- */

/* Core::buildDistanceField (core.cpp:33-37, core.hpp:37-59) from `map`; ctx remembers `map` as the
 * distance-field map used by rbv_ext_rot_vel / rbv_update_inverse_depth (core.hpp:39,63,78). */
int rbv_build_distance_field(rbv_ctx* ctx, rbv_map* map);

/* EdgeMap::rotateKeylines (edge_map.cpp:58-71); R: B row-major 3x3. */
int rbv_rotate_keylines(rbv_ctx* ctx, rbv_map* map, const float* R9);

/* EdgeMap::estimateQuantile (edge_map.cpp:39-56) */
int rbv_estimate_quantile(rbv_ctx* ctx, rbv_map* map, float percentile, int num_bins, float* out);

/* Core::tryVel (core.cpp:78-148): one energy/Jacobian evaluation at vel[B][3] with sigma_rho cut
 * sigma_rho_min[B]. Residuals live on the device with the map (reset with rbv_reset_residuals).
 * Outputs per lane: JtJ[9], JtF[3], score. */
int rbv_reset_residuals(rbv_ctx* ctx, rbv_map* map);
int rbv_try_vel(rbv_ctx* ctx, rbv_map* map, const float* vel3, const float* sigma_rho_min, float* JtJ9,
                float* JtF3, float* score);

/* Core::minimizeVel (core.cpp:150-189): quantile + (1+iterations) tryVel + 3x3 LM steps, all on device.
 * vel3 in/out [B][3], Rvel9 out [B][9], score out [B]. Any output pointer may be NULL (stays on device
 * for the following stages). */
int rbv_minimize_vel(rbv_ctx* ctx, rbv_map* old_map, float* vel3, float* Rvel9, float* score);

/* EdgeMap::forwardMatch (edge_map.cpp:73-99): old → new; n_matches[B] optional. */
int rbv_forward_match(rbv_ctx* ctx, rbv_map* old_map, rbv_map* new_map, int* n_matches);

/* Core::extRotVel (core.cpp:191-261) on the distance-field map. Wx36/X6 per lane, ok[B] = 1/0. */
int rbv_ext_rot_vel(rbv_ctx* ctx, rbv_map* new_map, const float* vel3, float* Wx36, float* X6, int* ok);

/* EdgeMap::directedMatch + searchMatch (edge_map.cpp:101-218): new searches in old. */
int rbv_directed_match(rbv_ctx* ctx, rbv_map* new_map, rbv_map* old_map, const float* vel3,
                       const float* Rvel9, const float* Rback9, float max_radius, int* kf_matches,
                       int* n_matches);

/* EdgeMap::regularize1Iter (edge_map.cpp:220-259) */
int rbv_regularize(rbv_ctx* ctx, rbv_map* map, int* n_regularized);

/* Core::updateInverseDepth (core.cpp:417-456) on the distance-field map */
int rbv_update_inverse_depth(rbv_ctx* ctx, const float* vel3);

/* ---- fused per-frame graphs (CUDA-graph captured; what Rebvio::stateEstimationProcess calls in order) --- */

/* G1 = buildDistanceField(new) ; rotateKeylines(old, R) ; minimizeVel(old) ; forwardMatch(old,new) ;
 *      extRotVel(new)   (rebvio.cpp:142,165,169,172,177). One launch of a captured graph + one D2H.
 * in : R9[B][9]      out: vel3[B][3] (Vg), Rvel9[B][9] (P_Vg), Wx36, X6, ok[B]                         */
int rbv_track_g1(rbv_ctx* ctx, rbv_map* old_map, rbv_map* new_map, const float* R9, float* vel3,
                 float* Rvel9, float* Wx36, float* X6, int* ok);

/* G2 = rotateKeylines(old, R0) ; directedMatch(new, old, V, P_V, Rgva) ; regularize1Iter(new) ;
 *      updateInverseDepth(V)   (rebvio.cpp:223/232,245,256,259). `min_matches`: lanes whose directed-match
 * count is below it skip regularise + Kalman exactly as rebvio.cpp:247-260 does.                        */
int rbv_track_g2(rbv_ctx* ctx, rbv_map* old_map, rbv_map* new_map, const float* R0_9, const float* V3,
                 const float* PV9, const float* Rgva9, float max_radius, int min_matches, int* n_matches,
                 int* kf_matches);

/* Host block between the two halves, visual-only subset of rebvio.cpp:183-233 (gyro prior and IMU filters switched off;
 * those stay in the reference's own host code): R0 = exp(X[3:6]) (rebvio.cpp:194-196),
 * V = R0 Vg + X[0:3], P_V = inv(Wx)[0:3][0:3] (rebvio.cpp:198-200), Rgva = (R0 R^T)^T. Pure host arithmetic in double,
 * no device access; all arrays [lanes][...]. A singular Wx gives P_V = 1e10 I.                                        */
int rbv_visual_fusion(int lanes, const float* R9, const float* Vg3, const float* Wx36, const float* X6, float* R0_9,
                      float* V3, float* PV9, float* Rgva9);

/* One tracker frame in one call: rbv_track_g1, rbv_visual_fusion, rbv_track_g2 (what a replay driver without IMU
 * filters does per frame). R9 = rotation of the camera since the previous frame, [lanes][9] row-major (the call
 * rotates the old map by its transpose, rebvio.cpp:165). Outputs (any may be NULL): Vg3, PVg9 (minimizeVel),
 * Wx36, X6, ok (extRotVel), V3 (fused velocity), n_matches / kf_matches (directedMatch), n_keylines_new /
 * auto_threshold_new (size and tuned threshold of new_map, so that rbv_detect* can be called without outputs and the
 * frame needs no device synchronisation of its own for them).                                                      */
int rbv_track(rbv_ctx* ctx, rbv_map* old_map, rbv_map* new_map, const float* R9, float max_radius, int min_matches,
              float* Vg3, float* PVg9, float* Wx36, float* X6, int* ok, float* V3, int* n_matches, int* kf_matches,
              int* n_keylines_new, float* auto_threshold_new);

/* ---- IMU fusion block on the device --------------------------------------------------------
 * The part of Rebvio::stateEstimationProcess between the tracker's two halves, per lane and device-resident:
 * gyro-bias initialisation and pre-rotation (rebvio.cpp:146-165), Core::gyroBiasCorrection (core.cpp:264-287),
 * Core::estimateLs4Acceleration / estimateMeanAcceleration (core.cpp:289-352), Core::estimateBias + SABEstimator
 * (core.cpp:355-414, sab_estimator.cpp:21-165), the composition of R, Rgva and V (rebvio.cpp:183-233) and the
 * incremental pose (rebvio.cpp:255-268). The IMU sample integrator (types::IntegratedImu, imu.hpp:33-150) stays on
 * the host: its result is the per-frame input below.                                                              */
typedef struct rbv_imu_cfg {            /* types::ImuStateConfig, imu.hpp:153-168 */
  float g_norm, g_uncertainty, g_norm_uncertainty, acc_std_dev, gyro_std_dev, gyro_bias_std_dev, vbias_std_dev;
  float scale_std_dev_mult, scale_std_dev_max, scale_stdd_dev_init;
  int init_bias, init_bias_frame_num;
  float init_bias_guess[3];
} rbv_imu_cfg;
void rbv_default_imu_cfg(rbv_imu_cfg* cfg);

typedef struct rbv_imu_in {             /* one lane, one frame: IntegratedImu::get() of the NEW map (rebvio.cpp:145) */
  float R[9];                           /* R()      integrated gyro rotation, row-major */
  float gyro[3];                        /* gyro()   mean rate */
  float cacc[3];                        /* cacc()   compensated mean acceleration */
  float dt_s;                           /* dt_s()   integration interval */
  float frame_dt;                       /* (new ts - old ts) / 1e6, rebvio.cpp:182 */
} rbv_imu_in;

typedef struct rbv_imu_out {            /* one lane, one frame */
  float Rt[9];                          /* R.T() handed to the first rotateKeylines (rebvio.cpp:163-165) */
  float R[9], R0[9], Rgva[9];           /* R after the visual correction; the rotation applied by the second
                                           rotateKeylines (R0 or R0gva); Rgva (rebvio.cpp:192-231) */
  float V[3], P_V[9];                   /* translation and covariance handed to directedMatch */
  float Xgv[6], Xgva[6];
  float Bg[3], Av[3], As[3], X[7];      /* gyro bias, visual / sensed acceleration, filter state after the frame */
  float orientation[3], position[3];    /* types::Odometry (rebvio.cpp:270-273) */
  float K;                              /* scale (rebvio.cpp:215) */
  int initialized, stopped, num_frames; /* stopped: 1 = NaN in V (rebvio.cpp:236-241) */
} rbv_imu_out;

/* (re)create the per-lane filter state = the state of a freshly constructed Rebvio (rebvio.hpp:92-101,
 * rebvio.cpp:95-115, sab_estimator.hpp:56-71). Must precede the calls below. */
int rbv_imu_reset(rbv_ctx* ctx, const rbv_imu_cfg* cfg);
/* stage call 1 (rebvio.cpp:146-165): bias initialisation, R = imu.R corrected by the bias. Rt9 [B][9] (may be NULL)
 * receives R.T(), the argument of the first rotateKeylines; it is also left in the device slot rbv_track_imu uses. */
int rbv_imu_prerotate(rbv_ctx* ctx, const rbv_imu_in* in /*[B]*/, float* Rt9);
/* stage call 2 (rebvio.cpp:183-233, 255-273): everything between extRotVel and the second rotateKeylines, plus
 * the pose. Inputs are the tracker's results [B][..]: Vg3 (minimizeVel), Wx36 / X6 (extRotVel). */
int rbv_imu_fuse(rbv_ctx* ctx, const float* Vg3, const float* Wx36, const float* X6, rbv_imu_out* out /*[B]*/);
/* One tracker frame with the IMU filters, kernels only: prerotate, first half (rbv_track_g1's sequence), fuse,
 * second half (rbv_track_g2's sequence), ONE device synchronisation at the end. Outputs as rbv_track (any may be
 * NULL) plus the fusion record per lane. Lanes whose directed-match count is below min_matches skip regularise and
 * Kalman update as in rbv_track_g2; the caller stops such a lane (rebvio.cpp:247-252).                            */
int rbv_track_imu(rbv_ctx* ctx, rbv_map* old_map, rbv_map* new_map, const rbv_imu_in* in, float max_radius,
                  int min_matches, float* Vg3, float* PVg9, float* Wx36, float* X6, int* ok, rbv_imu_out* out,
                  int* n_matches, int* kf_matches, int* n_keylines_new, float* auto_threshold_new);

/* ---- host mirror (EdgeMap::keylines(), operator[], mask(), size(), threshold(); edge_map.hpp:40-75) ---- */

int rbv_map_size(rbv_ctx* ctx, rbv_map* map, int* n_keylines /*[B]*/);
int rbv_map_threshold(rbv_ctx* ctx, rbv_map* map, float* threshold /*[B]*/);
int rbv_map_set_threshold(rbv_ctx* ctx, rbv_map* map, const float* threshold /*[B]*/);
/* AoS download of one lane: keylines (84 B each, capacity `cap`), detection pixel index per keyline
 * (= key of EdgeMap::mask()), returns count in *n. */
int rbv_map_download(rbv_ctx* ctx, rbv_map* map, int lane, rbv_keyline* keylines, uint32_t* pixel_idx,
                     int cap, int* n);
/* AoS upload of one lane (teacher forcing / callers that mutate keylines()). pixel_idx may be NULL to keep
 * the mask untouched. */
int rbv_map_upload(rbv_ctx* ctx, rbv_map* map, int lane, const rbv_keyline* keylines,
                   const uint32_t* pixel_idx, int n);

/* ---- debug taps used by the parity tests ---------------------------------------------------------------- */
/* which: 0 = scale0, 1 = scale1, 2 = DoG, 3 = squared gradient magnitude, 4 = the float frame the front end worked on
 * (after the x3 conversion / undistortion of the host-frame paths); out: rows*cols float32 */
int rbv_debug_read_image(rbv_ctx* ctx, int lane, int which, float* out);
/* id mask of EdgeDetector (keylines_mask_, edge_detector.hpp:86) for the last detect: rows*cols int32 */
int rbv_debug_read_mask(rbv_ctx* ctx, rbv_map* map, int lane, int* out);
/* distance field: id and distance per pixel (core.hpp:15-18); distance is only defined where id >= 0 */
int rbv_debug_read_field(rbv_ctx* ctx, int lane, int* id, int* distance);
int rbv_debug_read_residuals(rbv_ctx* ctx, rbv_map* map, int lane, float* out, int cap);
/* one box pass (createIntegralImage + average, scale_space.cpp:48-128) of width index `pass` of scale
 * `scale` applied to a host image; for the first-kernel parity test. */
int rbv_debug_box_pass(rbv_ctx* ctx, int scale, int pass, const float* in, float* out);
/* One box pass (createIntegralImage + average, scale_space.cpp:48-128: the arithmetic of rbv_debug_box_pass) of ONE image plane
 * that allocates nothing. NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it). d_in / d_out: DEVICE pointers to rows*cols
 * float32 (may not alias), on the CURRENT device; width: odd box width 3..15; cuda_stream: cudaStream_t (NULL = default stream).
 * Asynchronous: returns after the two kernel launches. No context is needed and no cudaMalloc / cudaFree / allocating routine is
 * called: the row-prefix scratch between the two kernels is a static device array of RBV_NOALLOC_MAX_PIXELS floats (4 MiB, held by
 * every device that loads the library until exit); rows*cols beyond that is rejected with RBV_ERR_ARG and a message. All calls
 * share that array (it keeps the last call's contents; nothing in it has to start at zero): calls that overlap on the GPU - two
 * streams without ordering between them - race. Calls on one stream are safe. */
#define RBV_NOALLOC_MAX_PIXELS (1024 * 1024)
int rbv_box_pass_noalloc(const float* d_in, float* d_out, int rows, int cols, int width, void* cuda_stream);
/* the 3x25 plane-fit pseudo inverse and box widths the context computed */
int rbv_debug_read_setup(rbv_ctx* ctx, float* pinv75, int* widths6);
/* LM trace of the last rbv_minimize_vel: per tryVel call t=0..iterations: vel[3], JtJ[9], JtF[3], score
 * (16 floats each); out capacity (iterations+1)*16 floats per lane. */
int rbv_debug_read_lm_trace(rbv_ctx* ctx, int lane, float* out, int cap_floats);
/* mean number of mask probes per keyline in the last directed match (P-bar of the byte model) */
int rbv_debug_read_probe_count(rbv_ctx* ctx, int lane, unsigned long long* probes);
/* debug: one accumulator's ORDERED-mode addends of one lane, keylines [0, n), after the last evaluation kernel (plane 0..9: tryVel in accumulator
   order, 100..126: extRotVel); plane -1: the chain kernel's running sum of accumulator 0 after every 256-keyline chunk (RBV_DEBUG_CHAIN=1) */
int rbv_debug_read_addends(rbv_ctx* ctx, int plane, int lane, float* out, int n);

#ifdef __cplusplus
}
#endif
#endif /* REBVIO_B200_H_ */
