// The reference's call shapes against the adapter, with the reference's own TooN types (built only where /root/reference/TooN is
// present - the build container; the binary travels to the GPU box): construction as in Rebvio::Rebvio (rebvio.cpp:19-30: EdgeDetector and
// Core each from their own COPY of the camera), detect(types::Image&)-shaped call, and the tracker half of
// Rebvio::stateEstimationProcess (rebvio.cpp:142-259) statement by statement with TooN vectors / matrices / SO3 / Cholesky, visual-only
// fusion branch (the IMU filters are the device block behind rbv_track_imu; see rebvio.hpp). The one source change a maintainer makes is
// the namespace alias below.
#include <TooN/TooN.h>
#include <TooN/Cholesky.h>
#include <TooN/so3.h>

#include <cstdio>
#include <vector>

#include "rebvio_b200/adapter.hpp"

namespace rebvio = rebvio_b200;
namespace types {   // types/definitions.hpp:17-34
using Float = float;
using Vector3f = TooN::Vector<3, Float>;
using Vector6f = TooN::Vector<6, Float>;
using Matrix3f = TooN::Matrix<3, 3, Float>;
using Matrix6f = TooN::Matrix<6, 6, Float>;
}  // namespace types

int main(int argc, char**) {
  rebvio::Camera camera_;
  rebvio::Camera::SharedPtr det_cam = std::make_shared<rebvio::Camera>(camera_), core_cam = std::make_shared<rebvio::Camera>(camera_);
  rebvio::EdgeDetectorConfig::SharedPtr det_cfg = std::make_shared<rebvio::EdgeDetectorConfig>();
  rebvio::CoreConfig::SharedPtr core_cfg = std::make_shared<rebvio::CoreConfig>();
  rebvio::EdgeDetector edge_detector_(det_cam, det_cfg);   // edge_detector.hpp:41
  rebvio::Core core_(core_cam, core_cfg);                  // core.hpp:99
  if (edge_detector_.device() != core_.device()) { std::puts("detector and core did not meet on one device"); return 2; }
  if (argc < 2) { std::puts("reference signatures compiled and linked"); return 0; }   // no GPU needed for the link check

  const int rows = camera_.rows_, cols = camera_.cols_;
  std::vector<float> px[2];
  rebvio::types::FloatImage img[2];
  for (int k = 0; k < 2; ++k) {
    px[k].resize((size_t)rows * cols);
    for (int r = 0; r < rows; ++r)
      for (int c = 0; c < cols; ++c) px[k][(size_t)r * cols + c] = (float)((((c + 2 * k) / 16 + r / 16) % 2) * 300 + 100);
    img[k].ts_us = 1000000 + 50000 * k;
    img[k].data.rows = rows; img[k].data.cols = cols; img[k].data.step = sizeof(float) * cols; img[k].data.data = px[k].data();
  }
  rebvio::EdgeMap::SharedPtr old_edge_map = edge_detector_.detect(img[0]);   // rebvio.cpp:70
  rebvio::EdgeMap::SharedPtr new_edge_map = edge_detector_.detect(img[1]);
  const float I9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  rebvio::types::Imu s;
  s.ts = 1000000;
  new_edge_map->imu().add(s, I9);                                            // edge_map.hpp:70, rebvio.cpp:81

  // ---- rebvio.cpp:142-259 ----
  core_.buildDistanceField(new_edge_map);                                    // :142
  types::Matrix3f R = TooN::Identity;
  old_edge_map->rotateKeylines(R.T());                                       // :165
  types::Vector3f Vg = TooN::Zeros;                                          // :168
  types::Matrix3f P_Vg;
  core_.minimizeVel(old_edge_map, Vg, P_Vg);                                 // :169
  old_edge_map->forwardMatch(new_edge_map);                                  // :172
  types::Matrix6f W_Xv;
  types::Vector6f Xv;
  core_.extRotVel(new_edge_map, Vg, W_Xv, Xv);                               // :177
  TooN::SO3<types::Float> R0(Xv.slice<3, 3>());                              // :229-232 (visual-only branch)
  types::Vector3f V = R0.get_matrix() * Vg + Xv.slice<0, 3>();
  types::Matrix6f iW = TooN::Cholesky<6, types::Float>(W_Xv).get_inverse();
  types::Matrix3f P_V = iW.slice<0, 0, 3, 3>();
  types::Matrix3f R0m = R0.get_matrix();
  types::Matrix3f Rgva = (R0m * R.T()).T();
  old_edge_map->rotateKeylines(R0m);                                         // :232
  int kf_matches = 0;
  const int klm_num = new_edge_map->directedMatch(old_edge_map, V, P_V, Rgva, kf_matches, 40.0f);   // :245
  new_edge_map->regularize1Iter();                                           // :256
  core_.updateInverseDepth(V);                                               // :259
  // in-place edits of keylines() go back with commit() (the device owns the data; edge_map.hpp:50 hands out a reference)
  std::vector<rebvio::KeyLine>& kl = new_edge_map->keylines();
  const float rho0 = kl.empty() ? 0.0f : kl[0].rho;
  std::printf("keylines %d directed %d kf %d Vg %g %g %g V %g %g %g rho0 %g imu_samples %u\n", new_edge_map->size(), klm_num, kf_matches, Vg[0], Vg[1],
              Vg[2], V[0], V[1], V[2], rho0, new_edge_map->imu().samples());
  return (new_edge_map->size() > 1000 && klm_num > 500) ? 0 : 1;
}
