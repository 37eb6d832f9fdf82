// Compile/link check of the C++ adapter against the C ABI with TooN-like stand-in types (no device is touched unless
// run with an argument). Built by tests/test_capi.py::test_cpp_adapter_compiles_and_links.
#include <cstdio>
#include <vector>

#include "rebvio_b200/adapter.hpp"

struct V3 { float d[3]{}; float& operator[](int i) { return d[i]; } const float& operator[](int i) const { return d[i]; } };
struct V6 { float d[6]{}; float& operator[](int i) { return d[i]; } const float& operator[](int i) const { return d[i]; } };
struct M3 { float d[9]{}; float& operator()(int i, int j) { return d[i * 3 + j]; } const float& operator()(int i, int j) const { return d[i * 3 + j]; } };
struct M6 { float d[36]{}; float& operator()(int i, int j) { return d[i * 6 + j]; } const float& operator()(int i, int j) const { return d[i * 6 + j]; } };

int main(int argc, char**) {
  using namespace rebvio_b200;
  if (argc < 2) { std::puts("adapter linked"); return 0; }   // no GPU needed for the link check
  Camera cam;
  EdgeDetectorConfig::SharedPtr dc = std::make_shared<EdgeDetectorConfig>();
  EdgeMapConfig mc;
  CoreConfig::SharedPtr cc = std::make_shared<CoreConfig>();
  auto dev = std::make_shared<Device>(cam, *dc, mc, *cc);
  EdgeDetector detector(dev, dc);
  Core core(dev, cc);
  std::vector<float> img((size_t)cam.rows_ * cam.cols_);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (float)(((i % cam.cols_) / 16 + (i / cam.cols_) / 16) % 2) * 300.0f + 100.0f;
  EdgeMap::SharedPtr old_map = detector.detect(img.data(), 0, 0), new_map = detector.detect(img.data(), 0, 50000);
  M3 R, Rvel; R(0, 0) = R(1, 1) = R(2, 2) = 1;
  V3 vel; M6 Wx; V6 X; int kf = 0;
  core.buildDistanceField(new_map);
  old_map->rotateKeylines(R);
  float score = core.minimizeVel(old_map, vel, Rvel);
  int fm = old_map->forwardMatch(new_map);
  bool ok = core.extRotVel(new_map, vel, Wx, X);
  old_map->rotateKeylines(R);
  int dm = new_map->directedMatch(old_map, vel, Rvel, R, kf, 40.0f);
  int rg = new_map->regularize1Iter();
  core.updateInverseDepth(vel);
  std::printf("keylines %d score %g fwd %d ok %d directed %d reg %d first rho %g\n", new_map->size(), score, fm, (int)ok, dm, rg,
              new_map->size() ? (*new_map)[0].rho : 0.0f);
  return 0;
}
