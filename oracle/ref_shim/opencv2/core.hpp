// Minimal stand-in for <opencv2/core.hpp>, used ONLY to compile the unmodified reference sources
// (/root/reference/rebvio/src/*.cpp) into oracle/_ref/ as the parity checker / CPU baseline.
// TEST INFRASTRUCTURE — never linked into the product library.
//
// The reference's hot path uses cv::Mat purely as a ref-counted 2-D buffer
// (scale_space.cpp:36-38,49,70,188-189; edge_detector.cpp:25,74-78,88,116,144-160; camera.hpp:39-40).
// Copies MUST be shallow: ScaleSpace::dog()/mag() return cv::Mat by value and callers keep row pointers
// into that temporary (scale_space.cpp:195-201, edge_detector.cpp:75-78).
#pragma once
#include <cstddef>
#include <cstring>
#include <memory>
#include <vector>

#define CV_32FC1 5
#define CV_32SC1 4
#define CV_8UC1 0
#define CV_64FC1 6

namespace cv {

struct Scalar {
  double v[4];
  Scalar(double a = 0) : v{a, 0, 0, 0} {}
};

inline size_t shim_elem_size(int type) {
  switch (type) { case CV_8UC1: return 1; case CV_64FC1: return 8; default: return 4; }
}

class Mat {
 public:
  int rows{0}, cols{0};
  unsigned char* data{nullptr};

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(int r, int c, int type, const Scalar& s) {
    create(r, c, type);
    const size_t n = size_t(r) * c;
    if (type == CV_32FC1) { float* p = reinterpret_cast<float*>(data); for (size_t i = 0; i < n; ++i) p[i] = float(s.v[0]); }
    else if (type == CV_32SC1) { int* p = reinterpret_cast<int*>(data); for (size_t i = 0; i < n; ++i) p[i] = int(s.v[0]); }
    else if (type == CV_64FC1) { double* p = reinterpret_cast<double*>(data); for (size_t i = 0; i < n; ++i) p[i] = s.v[0]; }
    else { std::memset(data, int(s.v[0]), n); }
  }
  // wrap external memory (no ownership) — used by the harness to hand frames in without a copy
  Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data(static_cast<unsigned char*>(ext)), type_(type) {}

  void create(int r, int c, int type) {
    rows = r; cols = c; type_ = type;
    store_ = std::shared_ptr<unsigned char>(new unsigned char[size_t(r) * c * shim_elem_size(type) + 64],
                                            std::default_delete<unsigned char[]>());
    data = store_.get();
  }
  int type() const { return type_; }
  bool empty() const { return data == nullptr; }

  template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data) + size_t(r) * cols; }
  template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data) + size_t(r) * cols; }
  template <typename T> T& at(int r, int c) { return ptr<T>(r)[c]; }
  template <typename T> const T& at(int r, int c) const { return ptr<T>(r)[c]; }
  template <typename T> T& at(int i) { return reinterpret_cast<T*>(data)[i]; }
  template <typename T> const T& at(int i) const { return reinterpret_cast<const T*>(data)[i]; }

  // only u8/f32 -> f32 with scale is needed (rebvio.cpp:43)
  void convertTo(Mat& out, int type, double alpha = 1.0, double beta = 0.0) const {
    Mat o(rows, cols, type);
    const size_t n = size_t(rows) * cols;
    float* d = reinterpret_cast<float*>(o.data);
    if (type_ == CV_8UC1) for (size_t i = 0; i < n; ++i) d[i] = float(data[i] * alpha + beta);
    else { const float* s = reinterpret_cast<const float*>(data); for (size_t i = 0; i < n; ++i) d[i] = float(s[i] * alpha + beta); }
    out = o;
  }
  Mat clone() const {
    Mat o(rows, cols, type_);
    std::memcpy(o.data, data, size_t(rows) * cols * shim_elem_size(type_));
    return o;
  }

 private:
  int type_{0};
  std::shared_ptr<unsigned char> store_;
};

template <typename T> struct shim_type_of;
template <> struct shim_type_of<float> { static const int value = CV_32FC1; };
template <> struct shim_type_of<double> { static const int value = CV_64FC1; };
template <> struct shim_type_of<int> { static const int value = CV_32SC1; };

// cv::Mat_<T>(r,c) << a, b, c ...   (camera.hpp:39-40)
template <typename T>
class Mat_ : public Mat {
 public:
  Mat_(int r, int c) : Mat(r, c, shim_type_of<T>::value) {}
  struct Filler {
    Mat_<T> m; size_t i;
    Filler operator,(double v) { m.template ptr<T>()[i] = T(v); return Filler{m, i + 1}; }
    operator Mat() const { return m; }
  };
  Filler operator<<(double v) { this->template ptr<T>()[0] = T(v); return Filler{*this, 1}; }
};

}  // namespace cv
