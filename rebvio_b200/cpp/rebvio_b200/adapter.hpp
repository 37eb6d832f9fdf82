// adapter.hpp — header-only C++ host side above the C ABI (include/rebvio_b200.h): the reference's class API for the
// hot path, same class / member names, argument meaning and return conventions, so that Rebvio's two worker loops
// (reference rebvio/src/rebvio.cpp:56-293) compile against it unchanged apart from the namespace:
//
//   rebvio::EdgeDetector::detect            edge_detector.hpp:50      -> rebvio_b200::EdgeDetector::detect
//   rebvio::Core::{buildDistanceField, minimizeVel, extRotVel, updateInverseDepth}   core.hpp:113,152,162,190
//   rebvio::EdgeMap::{rotateKeylines, forwardMatch, directedMatch, regularize1Iter, estimateQuantile, size, keylines,
//                     operator[], ts_us, threshold, mask}             edge_map.hpp:40-109
//
// Vector / matrix arguments are templates over anything indexable like TooN's types (v[i], M(i,j)): in the reference
// tree they are rebvio::types::Vector3f / Matrix3f / Matrix6f / Vector6f (types/definitions.hpp:20-34).
// The device keeps the keylines as SoA; `keylines()` / `operator[]` / `mask()` materialise the 84-byte AoS mirror lazily
// and `commit()` pushes host-side edits back (SURVEY.md 7 "Host<->device coherence at an AoS API").
// One context = one lane here (the reference's single stream); the batched multi-lane form is the C ABI itself.
// Reference signatures carried as they are: EdgeDetector(Camera::SharedPtr, cfg), Core(Camera::SharedPtr, cfg), detect(types::Image&)
// for a cv::Mat-shaped image, EdgeMap::imu(). tests/cpp/ref_signature_test.cpp compiles the tracker half of Rebvio's estimation loop
// (rebvio.cpp:142-259) against them with the reference's own TooN vector / matrix types.
// Errors: the reference's hot path throws nothing and reports through return values; C-ABI failures (no device, CUDA
// error) surface as std::runtime_error carrying rbv_last_error() — there is no CPU fallback to fall through to.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include <mutex>

#include "rebvio_b200.h"
#include "types.hpp"

namespace rebvio_b200 {

inline void check(int rc, const char* what) {
  if (rc != RBV_OK) throw std::runtime_error(std::string(what) + ": " + rbv_last_error());
}

struct Camera {  // fields of rebvio::Camera read by the hot path (camera.hpp:61-72)
  float fm_, cx_, cy_;
  int rows_, cols_;
  using SharedPtr = std::shared_ptr<Camera>;
  Camera() { rbv_camera c; rbv_default_camera(&c); fm_ = c.fm; cx_ = c.cx; cy_ = c.cy; rows_ = c.rows; cols_ = c.cols; }
};
struct EdgeDetectorConfig : rbv_detector_cfg {
  using SharedPtr = std::shared_ptr<EdgeDetectorConfig>;
  EdgeDetectorConfig() { rbv_default_detector_cfg(this); }
};
struct EdgeMapConfig : rbv_map_cfg {
  using SharedPtr = std::shared_ptr<EdgeMapConfig>;
  EdgeMapConfig() { rbv_default_map_cfg(this); }
};
struct CoreConfig : rbv_core_cfg {
  using SharedPtr = std::shared_ptr<CoreConfig>;
  CoreConfig() { rbv_default_core_cfg(this); }
};
using KeyLine = rbv_keyline;  // same 84-byte layout as rebvio::types::KeyLine (keyline.hpp:24-60)

// shared device context: EdgeDetector + Core of one pipeline talk to the same rbv_ctx
class Device {
 public:
  using SharedPtr = std::shared_ptr<Device>;
  Device(const Camera& cam, const rbv_detector_cfg& det, const rbv_map_cfg& map, const rbv_core_cfg& core, int device = 0,
         int reduce_mode = RBV_REDUCE_ORDERED)
      : cam_(cam), det_(det), map_(map), core_(core), device_(device), reduce_mode_(reduce_mode) { create(); }
  ~Device() { if (ctx_) rbv_destroy(ctx_); }
  Device(const Device&) = delete;
  Device& operator=(const Device&) = delete;
  rbv_ctx* ctx() { if (!ctx_) create(); return ctx_; }
  int keylines_max() const { return det_.keylines_max; }

  // ---- the reference's own constructor shapes: EdgeDetector(Camera::SharedPtr, cfg), Core(Camera::SharedPtr, cfg) (edge_detector.hpp:41,
  // core.hpp:99). The reference builds the two independently (rebvio.cpp:23-26: each from its own COPY of the camera); here both must
  // end up on one device context, created once both configurations are known. A detector constructed from a camera opens a pending
  // Device; the next Core constructed from a camera with the same intrinsics joins it (and vice versa); the context itself is created
  // at the first device call. Explicit sharing (the Device::SharedPtr constructors) is the unambiguous form when several pipelines
  // of one camera model live in a process.
  static SharedPtr forDetector(const Camera& cam, const rbv_detector_cfg& det) { return join(cam, &det, nullptr); }
  static SharedPtr forCore(const Camera& cam, const rbv_core_cfg& core) { return join(cam, nullptr, &core); }
  void setMapConfig(const rbv_map_cfg& m) { if (!ctx_) map_ = m; }

 private:
  Device(const Camera& cam) : cam_(cam), device_(0), reduce_mode_(RBV_REDUCE_ORDERED) {
    rbv_default_detector_cfg(&det_); rbv_default_map_cfg(&map_); rbv_default_core_cfg(&core_);
  }
  void create() {
    rbv_camera c{cam_.fm_, cam_.cx_, cam_.cy_, cam_.rows_, cam_.cols_};
    check(rbv_create(&c, &det_, &map_, &core_, device_, 1, &ctx_), "rbv_create");
    check(rbv_set_reduce_mode(ctx_, reduce_mode_), "rbv_set_reduce_mode");
  }
  static bool same(const Camera& a, const Camera& b) {
    return a.fm_ == b.fm_ && a.cx_ == b.cx_ && a.cy_ == b.cy_ && a.rows_ == b.rows_ && a.cols_ == b.cols_;
  }
  static SharedPtr join(const Camera& cam, const rbv_detector_cfg* det, const rbv_core_cfg* core) {
    static std::mutex mtx;
    static std::vector<std::weak_ptr<Device>> pending;
    std::lock_guard<std::mutex> lk(mtx);
    for (auto it = pending.begin(); it != pending.end();) {
      SharedPtr d = it->lock();
      if (!d || d->ctx_) { it = pending.erase(it); continue; }
      if (same(d->cam_, cam) && ((det && !d->has_det_) || (core && !d->has_core_))) {
        if (det) { d->det_ = *det; d->has_det_ = true; } else { d->core_ = *core; d->has_core_ = true; }
        if (d->has_det_ && d->has_core_) pending.erase(it);
        return d;
      }
      ++it;
    }
    SharedPtr d(new Device(cam));
    if (det) { d->det_ = *det; d->has_det_ = true; } else { d->core_ = *core; d->has_core_ = true; }
    pending.push_back(d);
    return d;
  }
  Camera cam_;
  rbv_detector_cfg det_;
  rbv_map_cfg map_;
  rbv_core_cfg core_;
  int device_, reduce_mode_;
  bool has_det_ = false, has_core_ = false;

 public:
  // the two worker loops of Rebvio get a CUDA stream each (RBV_TUNE_FE_STREAM): detect of frame n+1 runs next to the tracker of n-1, n
  // field_in_front_end: the distance field of a map is built behind its detect on that stream as well (RBV_TUNE_FIELD_IN_FE; pays with
  // many lanes, where the tracker is the longer chain - with one lane the two loops are balanced without it)
  void useFrontEndStream(bool on = true, bool field_in_front_end = false) {
    check(rbv_set_tuning(ctx_, RBV_TUNE_FE_STREAM, on ? 1 : 0), "rbv_set_tuning");
    check(rbv_set_tuning(ctx_, RBV_TUNE_FIELD_IN_FE, (on && field_in_front_end) ? 1 : 0), "rbv_set_tuning");
  }
  // Camera::undistort on the device (camera.hpp:54-58): K_ = [fm 0 cx; 0 fm cy; 0 0 1], D_ = (k1, k2, p1, p2, k3); raw frames from then on
  void setUndistort(float fm, float cx, float cy, float k1, float k2, float p1, float p2, float k3) {
    const float K4[4] = {fm, fm, cx, cy}, D5[5] = {k1, k2, p1, p2, k3};
    check(rbv_set_undistort(ctx_, K4, D5), "rbv_set_undistort");
  }

 private:
  rbv_ctx* ctx_ = nullptr;
};

class EdgeMap {
 public:
  using SharedPtr = std::shared_ptr<EdgeMap>;
  EdgeMap(Device::SharedPtr dev, uint64_t ts_us) : dev_(dev), ts_us_(ts_us) { check(rbv_map_create(dev_->ctx(), &map_), "rbv_map_create"); }
  ~EdgeMap() { rbv_map_release(dev_->ctx(), map_); }
  EdgeMap(const EdgeMap&) = delete;
  EdgeMap& operator=(const EdgeMap&) = delete;

  rbv_map* handle() const { return map_; }
  uint64_t ts_us() const { return ts_us_; }
  // EdgeMap::imu() (edge_map.hpp:70): the IMU samples integrated up to this map's frame (rebvio.cpp:77-84 adds them, :145 reads them)
  types::IntegratedImu& imu() { return imu_; }
  // a pooled map is reused for a later frame (the device slab and the captured graphs of its map pairs stay alive)
  void reset(uint64_t ts_us) { ts_us_ = ts_us; invalidate(); }
  int size() { int n = 0; check(rbv_map_size(dev_->ctx(), map_, &n), "rbv_map_size"); return n; }
  float threshold() { float t = 0; check(rbv_map_threshold(dev_->ctx(), map_, &t), "rbv_map_threshold"); return t; }
  void threshold(const float& t) { check(rbv_map_set_threshold(dev_->ctx(), map_, &t), "rbv_map_set_threshold"); }

  // lazy AoS mirror (edge_map.hpp:40-50,75). Any device-side stage invalidates it (invalidate() is called by the adapters).
  std::vector<KeyLine>& keylines() { pull(); return keylines_; }
  KeyLine& operator[](int idx) { pull(); return keylines_[idx]; }
  std::unordered_map<unsigned int, unsigned int>& mask() { pull(); return mask_; }
  void commit() {  // push host-side edits of keylines() back to the device
    if (!fresh_) return;
    check(rbv_map_upload(dev_->ctx(), map_, 0, keylines_.data(), pixels_.data(), (int)keylines_.size()), "rbv_map_upload");
  }
  void invalidate() { fresh_ = false; }

  // EdgeMap::estimateQuantile (edge_map.cpp:39-56)
  float estimateQuantile(float percentile, int num_bins) {
    float q = 0;
    check(rbv_estimate_quantile(dev_->ctx(), map_, percentile, num_bins, &q), "rbv_estimate_quantile");
    return q;
  }
  // EdgeMap::rotateKeylines (edge_map.cpp:58-71)
  template <class M3>
  void rotateKeylines(const M3& R) {
    float r[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r[i * 3 + j] = R(i, j);
    check(rbv_rotate_keylines(dev_->ctx(), map_, r), "rbv_rotate_keylines");
    invalidate();
  }
  // EdgeMap::forwardMatch (edge_map.cpp:73-99): this = old map, _map = new map; returns the match count
  int forwardMatch(SharedPtr map) {
    int n = 0;
    check(rbv_forward_match(dev_->ctx(), map_, map->map_, &n), "rbv_forward_match");
    map->invalidate();
    return n;
  }
  // EdgeMap::directedMatch (edge_map.cpp:186-218): this = new map searches in _map (old)
  template <class V3, class M3>
  int directedMatch(SharedPtr map, const V3& vel, const M3& Rvel, const M3& Rback, int& kf_matches, float max_radius) {
    float v[3], rv[9], rb[9];
    for (int i = 0; i < 3; ++i) { v[i] = vel[i]; for (int j = 0; j < 3; ++j) { rv[i * 3 + j] = Rvel(i, j); rb[i * 3 + j] = Rback(i, j); } }
    int n = 0, kf = 0;
    check(rbv_directed_match(dev_->ctx(), map_, map->map_, v, rv, rb, max_radius, &kf, &n), "rbv_directed_match");
    kf_matches = kf;
    invalidate();
    return n;
  }
  // EdgeMap::regularize1Iter (edge_map.cpp:220-259)
  int regularize1Iter() {
    int n = 0;
    check(rbv_regularize(dev_->ctx(), map_, &n), "rbv_regularize");
    invalidate();
    return n;
  }

 private:
  void pull() {
    if (fresh_) return;
    const int cap = dev_->keylines_max();
    keylines_.resize(cap);
    pixels_.resize(cap);
    int n = 0;
    check(rbv_map_download(dev_->ctx(), map_, 0, keylines_.data(), pixels_.data(), cap, &n), "rbv_map_download");
    keylines_.resize(n);
    pixels_.resize(n);
    mask_.clear();
    for (int i = 0; i < n; ++i) mask_.emplace(pixels_[i], (unsigned)i);   // edge_detector.cpp:112
    fresh_ = true;
  }
  Device::SharedPtr dev_;
  rbv_map* map_ = nullptr;
  uint64_t ts_us_;
  types::IntegratedImu imu_;
  bool fresh_ = false;
  std::vector<KeyLine> keylines_;
  std::vector<uint32_t> pixels_;
  std::unordered_map<unsigned int, unsigned int> mask_;
};

class EdgeDetector {
 public:
  EdgeDetector(Device::SharedPtr dev, EdgeDetectorConfig::SharedPtr config = std::make_shared<EdgeDetectorConfig>())
      : dev_(dev), config_(config) {}
  // the reference's constructor (edge_detector.hpp:41); see Device::forDetector for how detector and core meet on one context
  EdgeDetector(Camera::SharedPtr camera, EdgeDetectorConfig::SharedPtr config = std::make_shared<EdgeDetectorConfig>())
      : dev_(Device::forDetector(*camera, *config)), config_(config) {}
  EdgeDetectorConfig::SharedPtr config() { return config_; }
  Device::SharedPtr device() { return dev_; }
  // EdgeDetector::detect(types::Image&) (edge_detector.hpp:50): any image type shaped like the reference's types::Image, i.e. with
  // `ts_us` and a `data` member that looks like a CV_32FC1 cv::Mat (`ptr<float>(row)`, `step`): a real cv::Mat where OpenCV is
  // present, types::FloatMat otherwise. The frame is float32 0..765, already undistorted (rebvio.cpp:43-45).
  template <class ImageT>
  auto detect(ImageT& image) -> decltype((void)image.data.template ptr<float>(0), EdgeMap::SharedPtr()) {
    return detect(image.data.template ptr<float>(0), (size_t)image.data.step, (uint64_t)image.ts_us);
  }
  // EdgeDetector::detect (edge_detector.cpp:30-43). `image`: float32 rows x cols, values 0..765 (types::Image::data after
  // rebvio.cpp:43-45), row pitch in bytes (0 = dense). The servo mutates config_->threshold exactly as the reference does.
  EdgeMap::SharedPtr detect(const float* image, size_t pitch_bytes, uint64_t ts_us) {
    config_->threshold = rbv_servo_threshold(config_.get(), config_->threshold, keylines_count_);   // :33-36
    auto map = std::make_shared<EdgeMap>(dev_, ts_us);
    float at = 0;
    check(rbv_detect(dev_->ctx(), map->handle(), &image, pitch_bytes, &ts_us, &config_->threshold, &keylines_count_, &at), "rbv_detect");
    auto_threshold_ = at;
    return map;
  }
  // the same for the 8-bit frame the camera driver delivers: Rebvio::imageCallback's convertTo(x3) (rebvio.cpp:43) and, after
  // Device::setUndistort, Camera::undistort (rebvio.cpp:45) run on the device in front of the detector. `reuse`: a pooled map to
  // detect into instead of allocating one (the reference allocates a map per frame, edge_detector.cpp:31).
  EdgeMap::SharedPtr detect(const uint8_t* image, size_t pitch_bytes, uint64_t ts_us, EdgeMap::SharedPtr reuse = nullptr) {
    config_->threshold = rbv_servo_threshold(config_.get(), config_->threshold, keylines_count_);
    EdgeMap::SharedPtr map = reuse ? reuse : std::make_shared<EdgeMap>(dev_, ts_us);
    if (reuse) map->reset(ts_us);
    float at = 0;
    // enqueue without outputs, then wait for this detect alone: with Device::useFrontEndStream() the tracker stream keeps running
    check(rbv_detect_u8(dev_->ctx(), map->handle(), &image, pitch_bytes, &ts_us, &config_->threshold, nullptr, nullptr), "rbv_detect_u8");
    check(rbv_detect_wait(dev_->ctx(), map->handle(), &keylines_count_, &at, nullptr), "rbv_detect_wait");
    auto_threshold_ = at;
    return map;
  }
  int getNumKeylines() const { return keylines_count_; }

 private:
  Device::SharedPtr dev_;
  EdgeDetectorConfig::SharedPtr config_;
  int keylines_count_ = 0;
  float auto_threshold_ = 0;
};

class Core {
 public:
  Core(Device::SharedPtr dev, CoreConfig::SharedPtr config = std::make_shared<CoreConfig>()) : dev_(dev), config_(config) {}
  // the reference's constructor (core.hpp:99)
  Core(Camera::SharedPtr camera, CoreConfig::SharedPtr config = std::make_shared<CoreConfig>())
      : dev_(Device::forCore(*camera, *config)), config_(config) {}
  CoreConfig::SharedPtr config() { return config_; }
  Device::SharedPtr device() { return dev_; }
  // Core::buildDistanceField (core.cpp:33-37); keeps the map like DistanceField::map_ (core.hpp:39,63,78)
  void buildDistanceField(EdgeMap::SharedPtr map) {
    check(rbv_build_distance_field(dev_->ctx(), map->handle()), "rbv_build_distance_field");
    field_map_ = map;
  }
  // Core::minimizeVel (core.cpp:150-189): returns the score; vel in/out, Rvel out
  template <class V3, class M3>
  float minimizeVel(EdgeMap::SharedPtr map, V3& vel, M3& Rvel) {
    float v[3] = {vel[0], vel[1], vel[2]}, r[9], score = 0;
    check(rbv_minimize_vel(dev_->ctx(), map->handle(), v, r, &score), "rbv_minimize_vel");
    for (int i = 0; i < 3; ++i) { vel[i] = v[i]; for (int j = 0; j < 3; ++j) Rvel(i, j) = r[i * 3 + j]; }
    map->invalidate();
    return score;
  }
  // Core::extRotVel (core.cpp:191-261): false on NaN like the reference
  template <class V3, class M6, class V6>
  bool extRotVel(EdgeMap::SharedPtr map, const V3& vel, M6& Wx, V6& X) {
    float v[3] = {vel[0], vel[1], vel[2]}, w[36], x[6];
    int ok = 0;
    check(rbv_ext_rot_vel(dev_->ctx(), map->handle(), v, w, x, &ok), "rbv_ext_rot_vel");
    for (int i = 0; i < 6; ++i) { X[i] = x[i]; for (int j = 0; j < 6; ++j) Wx(i, j) = w[i * 6 + j]; }
    return ok != 0;
  }
  // Core::updateInverseDepth (core.cpp:417-422) on the distance-field map
  template <class V3>
  void updateInverseDepth(V3& vel) {
    float v[3] = {vel[0], vel[1], vel[2]};
    check(rbv_update_inverse_depth(dev_->ctx(), v), "rbv_update_inverse_depth");
    if (field_map_) field_map_->invalidate();
  }

 private:
  Device::SharedPtr dev_;
  CoreConfig::SharedPtr config_;
  EdgeMap::SharedPtr field_map_;
};

}  // namespace rebvio_b200
