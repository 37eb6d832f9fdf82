// rebvio.hpp — the pipeline façade above the C ABI: the reference's `rebvio::Rebvio` public interface
// (rebvio/include/rebvio/rebvio.hpp:38-113) with the same callbacks, the same two worker loops and the same hand-over
// of edge maps between them, but every per-frame computation is a device call:
//
//   Rebvio::imageCallback            rebvio.cpp:38-48    queues the 8-bit frame as delivered; x3 conversion and undistortion
//                                                        run on the device in front of the detector (rbv_detect_u8)
//   Rebvio::imuCallback              rebvio.cpp:50-54    queues the sample
//   Rebvio::dataAcquisitionProcess   rebvio.cpp:56-90    detect -> edge-image callbacks -> integrate the IMU samples with
//                                                        ts <= frame ts (IntegratedImu below) -> hand the map over
//   Rebvio::stateEstimationProcess   rebvio.cpp:92-293   ONE rbv_track_imu call per frame pair (pre-rotation, tracker first
//                                                        half, IMU fusion kernel, second half, pose) -> odometry callbacks
//   registerEdgeImageCallback / registerOdometryCallback  rebvio.cpp:301-307
//
// Differences from the reference, on purpose:
//  * the reference pushes the new edge map to the estimation queue BEFORE integrating its IMU samples (rebvio.cpp:72-80), so
//    the estimation thread can read a half-integrated record; here the map is handed over after the integration.
//  * edge maps come from a small pool (4 device slabs by default) instead of one allocation per frame; a map handed to an
//    edge-image callback is valid until the estimation thread has used it as the OLD map of the next pair.
//  * all filter state lives in the device context (no function-local statics, core.cpp:292-298,340-343), so any number of
//    Rebvio objects can share a process.
//  * the two loops have a CUDA stream each (RBV_TUNE_FE_STREAM) and rbv_track_imu drops the context lock while it waits for the device,
//    so detect of frame n+1 really runs next to the tracker of frames n-1, n, as on the reference's two CPU threads.
//  * waitIdle() lets a replay driver drain the queues deterministically; the reference has no such call (its test sleeps).
// Stop conditions are the reference's: NaN in the fused velocity or fewer directed matches than
// CoreConfig::global_min_matches_threshold end the run (rebvio.cpp:236-252); running() then returns false.
// There is no CPU fallback: construction throws std::runtime_error when the library finds no device.
#pragma once
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <queue>
#include <string>
#include <thread>
#include <vector>

#include "adapter.hpp"

namespace rebvio_b200 {

struct RebvioConfig {                 // rebvio::RebvioConfig (rebvio.hpp:29-33) + what the reference hard-codes in Camera
  EdgeDetectorConfig edge_detector;
  CoreConfig core;
  types::ImuStateConfig imu_state;
  EdgeMapConfig edge_map;
  Camera camera;                      // camera.hpp:26-37 defaults (EuRoC cam0)
  // camera.hpp:41-45: camera -> IMU extrinsics (EuRoC cam0), row-major
  float R_c2i[9] = {0.0148655429818f, -0.999880929698f, 0.00414029679422f, 0.999557249008f, 0.0149672133247f, 0.025715529948f,
                    -0.0257744366974f, 0.00375618835797f, 0.999660727178f};
  float t_c2i[3] = {-0.0216401454975f, -0.064676986768f, 0.00981073058949f};
  bool undistort = false;             // Camera::undistort on the device (rbv_set_undistort) with the coefficients below
  float distortion[5] = {-0.28340811f, 0.07395907f, 0.00019359f, 1.76187114e-05f, 0.0f};   // camera.hpp:35 (k1 k2 p1 p2 k3)
  int device = 0;
  int map_pool = 4;                   // >= 3: old + new on the tracker, one in the detector; a fourth lets the detector run a frame ahead
  bool two_streams = true;            // the acquisition loop's detect on the front-end stream, next to the estimation loop's tracker call
};

class Rebvio {
 public:
  explicit Rebvio(RebvioConfig& config)
      : config_(config), run_(true), stop_(false), num_frames_(0), busy_(0) {
    det_cfg_ = std::make_shared<EdgeDetectorConfig>(config_.edge_detector);
    core_cfg_ = std::make_shared<CoreConfig>(config_.core);
    dev_ = std::make_shared<Device>(config_.camera, *det_cfg_, config_.edge_map, *core_cfg_, config_.device);
    if (config_.undistort)
      dev_->setUndistort(config_.camera.fm_, config_.camera.cx_, config_.camera.cy_, config_.distortion[0], config_.distortion[1],
                         config_.distortion[2], config_.distortion[3], config_.distortion[4]);
    dev_->useFrontEndStream(config_.two_streams);
    check(rbv_imu_reset(dev_->ctx(), &config_.imu_state), "rbv_imu_reset");
    edge_detector_.reset(new EdgeDetector(dev_, det_cfg_));
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Rt_c2i_[i * 3 + j] = config_.R_c2i[j * 3 + i];
    const int pool = config_.map_pool < 3 ? 3 : config_.map_pool;
    for (int i = 0; i < pool; ++i) free_maps_.push_back(std::make_shared<EdgeMap>(dev_, 0));
    data_acquisition_thread_ = std::thread(&Rebvio::dataAcquisitionProcess, this);
    state_estimation_thread_ = std::thread(&Rebvio::stateEstimationProcess, this);
  }
  ~Rebvio() {
    stop_ = true;
    cv_.notify_all();
    data_acquisition_thread_.join();
    state_estimation_thread_.join();
    std::queue<Frame>().swap(edge_map_buffer_);   // maps release their device slabs before the context goes
    free_maps_.clear();
    held_old_.reset();
    edge_detector_.reset();
  }
  Rebvio(const Rebvio&) = delete;
  Rebvio& operator=(const Rebvio&) = delete;

  // rebvio.cpp:38-48. The frame is queued as delivered (8 bit); conversion and undistortion happen on the device.
  void imageCallback(types::Image&& image) {
    { std::lock_guard<std::mutex> g(mutex_); ++busy_; image_buffer_.push(std::move(image)); }
    cv_.notify_all();
  }
  // rebvio.cpp:50-54
  void imuCallback(types::Imu&& imu) {
    std::lock_guard<std::mutex> g(mutex_);
    imu_buffer_.push(imu);
  }
  // rebvio.cpp:301-307. Register before the first frame is delivered.
  void registerEdgeImageCallback(std::function<void(const types::Image&, EdgeMap::SharedPtr&)> cb) { edge_image_callbacks_.push_back(cb); }
  void registerOdometryCallback(std::function<void(types::Odometry&)> cb) { odometry_callbacks_.push_back(cb); }

  bool running() const { return run_; }                 // false once a stop condition of rebvio.cpp:236-252 has fired
  unsigned int numFrames() const { return num_frames_; }
  const rbv_imu_out& lastState() const { return last_; }   // filter state after the most recent frame pair (K, biases, ...)
  int lastMatches() const { return last_matches_; }
  // block until every queued image has gone through both loops (or the run has stopped)
  void waitIdle() {
    std::unique_lock<std::mutex> lk(mutex_);
    cv_idle_.wait(lk, [this] { return busy_ == 0 || !run_ || stop_; });
  }

 private:
  struct Frame { EdgeMap::SharedPtr map; rbv_imu_in imu; bool imu_ok; uint64_t ts_us; };

  void dataAcquisitionProcess() {   // rebvio.cpp:56-90
    try {
    for (;;) {
      types::Image img;
      EdgeMap::SharedPtr slot;
      {
        std::unique_lock<std::mutex> lk(mutex_);
        cv_.wait(lk, [this] { return stop_ || !run_ || (!image_buffer_.empty() && !free_maps_.empty()); });
        if (stop_ || !run_) break;
        img = std::move(image_buffer_.front());
        image_buffer_.pop();
        slot = free_maps_.front();
        free_maps_.pop_front();
      }
      EdgeMap::SharedPtr edge_map = edge_detector_->detect(img.data.data(), (size_t)img.cols, img.ts_us, slot);   // :70
      for (auto& cb : edge_image_callbacks_) cb(img, edge_map);                                                  // :71
      Frame f;
      f.map = edge_map; f.ts_us = img.ts_us; f.imu_ok = false;
      std::memset(&f.imu, 0, sizeof f.imu);
      {
        std::lock_guard<std::mutex> g(mutex_);
        while (!imu_buffer_.empty() && imu_buffer_.front().ts <= img.ts_us) {   // :77-84
          integrator_.add(imu_buffer_.front(), Rt_c2i_);
          imu_buffer_.pop();
        }
        // (new ts - old ts) / 1e6 of rebvio.cpp:182, evaluated where both stamps are at hand
        const float frame_dt = have_prev_ts_ ? (float)((double)(float)(img.ts_us - prev_ts_) / 1000000.0) : 0.0f;
        f.imu_ok = integrator_.get(Rt_c2i_, config_.t_c2i, frame_dt, &f.imu);
        if (!f.imu_ok) integrator_ = types::IntegratedImu();
        prev_ts_ = img.ts_us; have_prev_ts_ = true;
        edge_map_buffer_.push(f);
      }
      cv_.notify_all();
    }
    } catch (const std::exception& e) { fail(e.what()); }
  }

  void stateEstimationProcess() {   // rebvio.cpp:92-293
    try {
    for (;;) {
      Frame old_f, new_f;
      {
        std::unique_lock<std::mutex> lk(mutex_);
        cv_.wait(lk, [this] { return stop_ || !run_ || !edge_map_buffer_.empty(); });
        if (stop_ || !run_) break;
        if (!held_old_) {   // the very first map only ever serves as the old one (rebvio.cpp:124-128 waits for two)
          first_ = edge_map_buffer_.front(); held_old_ = first_.map; edge_map_buffer_.pop(); done_one();
          continue;
        }
        old_f = first_;
        new_f = edge_map_buffer_.front();
        edge_map_buffer_.pop();
      }
      bool stop_run = false;
      if (!new_f.imu_ok) {
        // fewer than two IMU samples in the interval: the reference would divide by zero (imu.hpp:81); stop the run
        std::fprintf(stderr, "rebvio_b200::Rebvio: fewer than two IMU samples before frame %llu, stopping\n", (unsigned long long)new_f.ts_us);
        stop_run = true;
      } else {
        int n_matches = 0, kf_matches = 0, ok = 0;
        rbv_imu_out out;
        check(rbv_track_imu(dev_->ctx(), old_f.map->handle(), new_f.map->handle(), &new_f.imu, core_cfg_->search_range /* rebvio.cpp:245 */,
                            (int)core_cfg_->global_min_matches_threshold, nullptr, nullptr, nullptr, nullptr, &ok, &out, &n_matches,
                            &kf_matches, nullptr, nullptr), "rbv_track_imu");
        old_f.map->invalidate();
        new_f.map->invalidate();
        last_ = out; last_matches_ = n_matches;
        types::Odometry odometry;                                                   // :270-273
        odometry.ts_us = new_f.ts_us;
        for (int i = 0; i < 3; ++i) { odometry.orientation[i] = out.orientation[i]; odometry.position[i] = out.position[i]; }
        for (auto& cb : odometry_callbacks_) cb(odometry);
        ++num_frames_;
        if (out.stopped || n_matches < (int)core_cfg_->global_min_matches_threshold) {   // :236-252
          std::fprintf(stderr, "rebvio_b200::Rebvio: %s at frame %llu, stopping\n", out.stopped ? "NaN in the velocity estimate" : "too few matches",
                       (unsigned long long)new_f.ts_us);
          stop_run = true;
        }
      }
      {
        std::lock_guard<std::mutex> g(mutex_);
        free_maps_.push_back(old_f.map);     // the old map's slab goes back to the detector
        first_ = new_f; held_old_ = new_f.map;
        done_one();
        if (stop_run) run_ = false;
      }
      cv_.notify_all();
      cv_idle_.notify_all();
      if (stop_run) break;
    }
    } catch (const std::exception& e) { fail(e.what()); }
  }

  void done_one() { if (busy_ > 0) --busy_; if (busy_ == 0) cv_idle_.notify_all(); }   // caller holds mutex_
  void fail(const char* what) {
    std::fprintf(stderr, "rebvio_b200::Rebvio: %s\n", what);
    { std::lock_guard<std::mutex> g(mutex_); run_ = false; }
    cv_.notify_all();
    cv_idle_.notify_all();
  }

  RebvioConfig config_;
  std::atomic<bool> run_, stop_;
  std::atomic<unsigned int> num_frames_;
  EdgeDetectorConfig::SharedPtr det_cfg_;
  CoreConfig::SharedPtr core_cfg_;
  Device::SharedPtr dev_;
  std::unique_ptr<EdgeDetector> edge_detector_;
  float Rt_c2i_[9];
  types::IntegratedImu integrator_;
  uint64_t prev_ts_ = 0;
  bool have_prev_ts_ = false;
  rbv_imu_out last_{};
  int last_matches_ = 0;

  std::mutex mutex_;                              // one lock for the three queues and the map pool
  std::condition_variable cv_, cv_idle_;
  int busy_;                                      // images delivered and not yet through the estimation loop
  std::queue<types::Image> image_buffer_;         // rebvio.hpp:102
  std::queue<types::Imu> imu_buffer_;             // rebvio.hpp:104
  std::queue<Frame> edge_map_buffer_;             // rebvio.hpp:106
  std::deque<EdgeMap::SharedPtr> free_maps_;
  Frame first_{};
  EdgeMap::SharedPtr held_old_;

  std::vector<std::function<void(const types::Image&, EdgeMap::SharedPtr&)>> edge_image_callbacks_;   // rebvio.hpp:109
  std::vector<std::function<void(types::Odometry&)>> odometry_callbacks_;                             // rebvio.hpp:110
  std::thread data_acquisition_thread_, state_estimation_thread_;                                    // rebvio.hpp:93-94
};

}  // namespace rebvio_b200
