// rebvio_replay — headless replay through the C++ pipeline façade (rebvio_b200/rebvio.hpp): camera frames + IMU samples from
// a file in, the reference's odometry log out. Stands in for the reference's ROS node (ros_rebvio/src/ros_rebvio.cpp:21-50
// forwards the topics to Rebvio::imageCallback / imuCallback, :89-126 publishes the odometry) and delivers the data in the
// deterministic order of the reference's replay test (ros_rebvio/test/test_ros_rebvio.cpp:29-45: every IMU sample up to an
// image's time stamp before that image).
//
//   rebvio_replay <sequence.rbvseq> <odometry.txt> [--imu-only] [--no-wait]
//
// sequence file (little endian, written by rebvio_b200/replay.py::write_sequence):
//   char magic[8] = "RBVSEQ1\0"; int32 rows, cols, n_frames, n_imu, keylines_ref, keylines_max; float32 fm, cx, cy;
//   uint64 ts_us[n_frames]; uint8 frames[n_frames][rows][cols]; uint64 imu_ts_us[n_imu]; float32 gyro[n_imu][3]; float32 acc[n_imu][3]
// --imu-only: no device; print the integrated IMU record of every frame interval (hex floats) — the CPU test of IntegratedImu.
// --no-wait : deliver everything at once (both worker loops run concurrently) instead of draining after every frame.
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <vector>

#include "rebvio_b200/rebvio.hpp"

using namespace rebvio_b200;

struct Sequence {
  int32_t rows = 0, cols = 0, n_frames = 0, n_imu = 0, keylines_ref = 0, keylines_max = 0;
  float fm = 0, cx = 0, cy = 0;
  std::vector<uint64_t> ts, imu_ts;
  std::vector<uint8_t> frames;
  std::vector<float> gyro, acc;
};

static bool load(const char* path, Sequence& s) {
  std::ifstream f(path, std::ios::binary);
  char magic[8];
  if (!f.read(magic, 8) || std::memcmp(magic, "RBVSEQ1", 8) != 0) return false;
  f.read((char*)&s.rows, 4); f.read((char*)&s.cols, 4); f.read((char*)&s.n_frames, 4); f.read((char*)&s.n_imu, 4);
  f.read((char*)&s.keylines_ref, 4); f.read((char*)&s.keylines_max, 4);
  f.read((char*)&s.fm, 4); f.read((char*)&s.cx, 4); f.read((char*)&s.cy, 4);
  if (!f || s.rows <= 0 || s.cols <= 0 || s.n_frames < 0 || s.n_imu < 0) return false;
  s.ts.resize(s.n_frames); s.imu_ts.resize(s.n_imu);
  s.frames.resize((size_t)s.n_frames * s.rows * s.cols);
  s.gyro.resize((size_t)s.n_imu * 3); s.acc.resize((size_t)s.n_imu * 3);
  f.read((char*)s.ts.data(), s.ts.size() * 8);
  f.read((char*)s.frames.data(), s.frames.size());
  f.read((char*)s.imu_ts.data(), s.imu_ts.size() * 8);
  f.read((char*)s.gyro.data(), s.gyro.size() * 4);
  f.read((char*)s.acc.data(), s.acc.size() * 4);
  return (bool)f;
}

static types::Imu sample(const Sequence& s, int j) {
  types::Imu m;
  m.ts = s.imu_ts[j];
  for (int i = 0; i < 3; ++i) { m.gyro[i] = s.gyro[(size_t)j * 3 + i]; m.acc[i] = s.acc[(size_t)j * 3 + i]; }
  return m;
}

int main(int argc, char** argv) {
  if (argc < 3) { std::fprintf(stderr, "usage: %s <sequence.rbvseq> <odometry.txt> [--imu-only] [--no-wait]\n", argv[0]); return 2; }
  bool imu_only = false, wait = true;
  for (int i = 3; i < argc; ++i) { if (!std::strcmp(argv[i], "--imu-only")) imu_only = true; else if (!std::strcmp(argv[i], "--no-wait")) wait = false; }
  Sequence s;
  if (!load(argv[1], s)) { std::fprintf(stderr, "cannot read %s\n", argv[1]); return 2; }
  FILE* out = std::fopen(argv[2], "w");
  if (!out) { std::fprintf(stderr, "cannot write %s\n", argv[2]); return 2; }
  RebvioConfig cfg;
  cfg.camera.rows_ = s.rows; cfg.camera.cols_ = s.cols; cfg.camera.fm_ = s.fm; cfg.camera.cx_ = s.cx; cfg.camera.cy_ = s.cy;
  if (s.keylines_ref > 0) cfg.edge_detector.keylines_ref = s.keylines_ref;
  if (s.keylines_max > 0) cfg.edge_detector.keylines_max = s.keylines_max;

  if (imu_only) {   // host integrator alone, in the order dataAcquisitionProcess uses it
    float Rt[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Rt[i * 3 + j] = cfg.R_c2i[j * 3 + i];
    types::IntegratedImu integ;
    int j = 0;
    for (int k = 0; k < s.n_frames; ++k) {
      while (j < s.n_imu && s.imu_ts[j] <= s.ts[k]) { integ.add(sample(s, j), Rt); ++j; }
      rbv_imu_in rec;
      std::memset(&rec, 0, sizeof rec);
      const float fdt = k ? (float)((double)(float)(s.ts[k] - s.ts[k - 1]) / 1000000.0) : 0.0f;
      if (!integ.get(Rt, cfg.t_c2i, fdt, &rec)) { integ = types::IntegratedImu(); std::fprintf(out, "%" PRIu64 " none\n", s.ts[k]); continue; }
      std::fprintf(out, "%" PRIu64, s.ts[k]);
      const float* p = (const float*)&rec;
      for (size_t i = 0; i < sizeof(rec) / 4; ++i) { uint32_t u; std::memcpy(&u, p + i, 4); std::fprintf(out, " %08x", u); }
      std::fprintf(out, "\n");
    }
    std::fclose(out);
    return 0;
  }

  int lines = 0, edge_images = 0;
  try {
    Rebvio rebvio(cfg);
    rebvio.registerOdometryCallback([&](types::Odometry& o) { std::fprintf(out, "%s\n", types::odometry_line(o).c_str()); ++lines; });
    rebvio.registerEdgeImageCallback([&](const types::Image&, EdgeMap::SharedPtr& m) { if (m->ts_us() || edge_images == 0) ++edge_images; });
    int j = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (int k = 0; k < s.n_frames && rebvio.running(); ++k) {
      while (j < s.n_imu && s.imu_ts[j] <= s.ts[k]) { rebvio.imuCallback(sample(s, j)); ++j; }
      types::Image img;
      img.ts_us = s.ts[k]; img.rows = s.rows; img.cols = s.cols;
      const size_t n = (size_t)s.rows * s.cols;
      img.data.assign(s.frames.begin() + k * n, s.frames.begin() + (k + 1) * n);
      rebvio.imageCallback(std::move(img));
      if (wait) rebvio.waitIdle();
    }
    rebvio.waitIdle();
    const double wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::fprintf(stderr, "rebvio_replay: %d frames in, %d edge images, %d odometry lines, scale K %.6f, last matches %d, running %d, %.1f frames/s wall\n",
                 s.n_frames, edge_images, lines, rebvio.lastState().K, rebvio.lastMatches(), (int)rebvio.running(), s.n_frames / wall_s);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "rebvio_replay: %s\n", e.what());
    std::fclose(out);
    return 1;
  }
  std::fclose(out);
  return 0;
}
