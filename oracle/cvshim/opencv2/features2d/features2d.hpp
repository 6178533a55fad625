// TEST INFRASTRUCTURE ONLY (oracle). Minimal stand-in for the OpenCV C++ API
// that the reference's external/brisk sources touch, so that those sources
// compile UNCHANGED from /root/reference into oracle/_ref/ (see oracle/Makefile).
// OpenCV C++ headers are not installed in this image; only the Python cv2 4.13
// wheel is, and the OpenCV-defined arithmetic below is pinned against it by
// tests/test_oracle_cv_pins.py (goldens in tests/golden/cv_pins.npz):
//   * Mat * Mat (cv::gemm small-matrix path: double accumulation, cast to float)
//   * cv::normalize(Vec3f) (norm accumulated in double, scale in double)
//   * cv::eigen on a 2x2 float Mat (Jacobi, reads the upper off-diagonal only,
//     eigenvalues sorted descending)
// Nothing under okvis2-x_b200/ (the product) includes this file.
#ifndef ORACLE_CVSHIM_FEATURES2D_HPP_
#define ORACLE_CVSHIM_FEATURES2D_HPP_

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <iostream>
#include <limits>
#include <string>
#include <vector>

#define CV_MAJOR_VERSION 4
#define CV_MINOR_VERSION 13

#define CV_CN_SHIFT 3
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAKETYPE(depth, cn) (CV_MAT_DEPTH(depth) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_16UC1 CV_MAKETYPE(CV_16U, 1)
#define CV_16SC1 CV_MAKETYPE(CV_16S, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC3 CV_MAKETYPE(CV_32F, 3)
#define CV_32FC6 CV_MAKETYPE(CV_32F, 6)

typedef unsigned char uchar;

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;

// Fixed vector with the float semantics of cv::Matx/cv::Vec (matx.hpp):
// element-wise ops in _Tp, dot accumulates in _Tp in index order.
template <typename T, int N>
struct Vec {
  T val[N];
  Vec() { for (int i = 0; i < N; ++i) val[i] = T(0); }
  Vec(T a, T b) { static_assert(N == 2, ""); val[0] = a; val[1] = b; }
  Vec(T a, T b, T c) { static_assert(N == 3, ""); val[0] = a; val[1] = b; val[2] = c; }
  T& operator[](int i) { return val[i]; }
  const T& operator[](int i) const { return val[i]; }
  T dot(const Vec& o) const {
    T s = 0;
    for (int i = 0; i < N; ++i) s += val[i] * o.val[i];
    return s;
  }
  Vec cross(const Vec& v) const {
    static_assert(N == 3, "");
    return Vec(val[1] * v.val[2] - val[2] * v.val[1],
               val[2] * v.val[0] - val[0] * v.val[2],
               val[0] * v.val[1] - val[1] * v.val[0]);
  }
};
template <typename T, int N>
inline Vec<T, N> operator+(const Vec<T, N>& a, const Vec<T, N>& b) {
  Vec<T, N> r;
  for (int i = 0; i < N; ++i) r.val[i] = a.val[i] + b.val[i];
  return r;
}
template <typename T, int N>
inline Vec<T, N> operator*(float alpha, const Vec<T, N>& a) {
  Vec<T, N> r;
  for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] * alpha);
  return r;
}
template <typename T, int N>
inline Vec<T, N> operator*(const Vec<T, N>& a, double alpha) {
  Vec<T, N> r;
  for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] * alpha);
  return r;
}
typedef Vec<float, 2> Vec2f;
typedef Vec<float, 3> Vec3f;
typedef Vec<float, 6> Vec6f;

// cv::normalize(Vec) (matx.hpp): nv = sqrt(sum of squares in double);
// v * (nv ? 1./nv : 0.) with the product in double, cast back.
template <typename T, int N>
inline Vec<T, N> normalize(const Vec<T, N>& v) {
  double s = 0;
  for (int i = 0; i < N; ++i) {
    double e = v.val[i];
    s += e * e;
  }
  double nv = std::sqrt(s);
  return v * (nv ? 1. / nv : 0.);
}

struct KeyPoint {
  Point2f pt;
  float size;
  float angle;
  float response;
  int octave;
  int class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct MatStep {
  size_t p[2];
  MatStep() { p[0] = p[1] = 0; }
  operator size_t() const { return p[0]; }
  size_t operator[](int i) const { return p[i]; }
  size_t& operator[](int i) { return p[i]; }
};

class Mat {
 public:
  int rows, cols;
  uchar* data;
  MatStep step;

  Mat() : rows(0), cols(0), data(nullptr), type_(0) {}
  Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), type_(0) { create(r, c, type); }
  // External, non-owning data (as cv::Mat(rows, cols, type, void*)).
  Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data(static_cast<uchar*>(ext)), type_(type) {
    step.p[1] = elemSize();
    step.p[0] = step.p[1] * static_cast<size_t>(c);
  }
  void create(int r, int c, int type) {
    if (data && r == rows && c == cols && type == type_) return;
    rows = r; cols = c; type_ = type;
    step.p[1] = elemSize();
    step.p[0] = step.p[1] * static_cast<size_t>(c);
    size_t bytes = step.p[0] * static_cast<size_t>(r);
    // Like cv::fastMalloc: uninitialised, 64-byte aligned.
    void* ptr = nullptr;
    if (posix_memalign(&ptr, 64, bytes ? bytes + 64 : 64) != 0) ptr = nullptr;
    store_.reset(static_cast<uchar*>(ptr), free);
    data = store_.get();
  }
  static Mat zeros(int r, int c, int type) {
    Mat m(r, c, type);
    std::memset(m.data, 0, m.step.p[0] * static_cast<size_t>(r));
    return m;
  }
  int type() const { return type_; }
  int depth() const { return CV_MAT_DEPTH(type_); }
  int channels() const { return (type_ >> CV_CN_SHIFT) + 1; }
  size_t elemSize1() const {
    static const size_t s[] = {1, 1, 2, 2, 4, 4, 8, 2};
    return s[depth()];
  }
  size_t elemSize() const { return elemSize1() * static_cast<size_t>(channels()); }
  bool isContinuous() const { return true; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  template <typename T> T& at(int r, int c) {
    return *reinterpret_cast<T*>(data + step.p[0] * static_cast<size_t>(r) + sizeof(T) * static_cast<size_t>(c));
  }
  template <typename T> const T& at(int r, int c) const {
    return *reinterpret_cast<const T*>(data + step.p[0] * static_cast<size_t>(r) + sizeof(T) * static_cast<size_t>(c));
  }
  template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + step.p[0] * static_cast<size_t>(r)); }

 private:
  int type_;
  std::shared_ptr<uchar> store_;
};

// Mat * Mat for CV_32FC1: cv::gemm generic path for small float matrices
// (GEMMSingleMul<float,double>): each output is sum_k double(a)*double(b),
// accumulated in double in k order, times alpha (=1), cast to float.
inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32FC1 && b.type() == CV_32FC1 && a.cols == b.rows);
  Mat d(a.rows, b.cols, CV_32FC1);
  for (int i = 0; i < a.rows; ++i)
    for (int j = 0; j < b.cols; ++j) {
      double s = 0;
      for (int k = 0; k < a.cols; ++k)
        s += static_cast<double>(a.at<float>(i, k)) * static_cast<double>(b.at<float>(k, j));
      d.at<float>(i, j) = static_cast<float>(s * 1.0);
    }
  return d;
}

namespace shim_detail {
inline float jhypot(float a, float b) {
  a = std::abs(a);
  b = std::abs(b);
  if (a > b) {
    b /= a;
    return a * std::sqrt(1 + b * b);
  }
  if (b > 0) {
    a /= b;
    return b * std::sqrt(1 + a * a);
  }
  return 0;
}
}  // namespace shim_detail

// cv::eigen for a 2x2 CV_32FC1 matrix without HAVE_EIGEN: JacobiImpl_<float>
// (modules/core/src/lapack.cpp) specialised to n = 2. Only A(0,1) is read as
// the off-diagonal; eigenvalues are returned in descending order.
inline bool eigen(const Mat& src, Vec2f& eigenvalues) {
  assert(src.type() == CV_32FC1 && src.rows == 2 && src.cols == 2);
  float W0 = src.at<float>(0, 0), W1 = src.at<float>(1, 1);
  float p = src.at<float>(0, 1);
  const float eps = FLT_EPSILON;
  if (std::abs(p) > eps) {
    float y = (W1 - W0) * 0.5f;
    float t = std::abs(y) + shim_detail::jhypot(p, y);
    float s = shim_detail::jhypot(p, t);
    (void)s;
    t = (p / t) * p;
    if (y < 0) t = -t;
    W0 -= t;
    W1 += t;
  }
  if (W0 < W1) std::swap(W0, W1);
  eigenvalues[0] = W0;
  eigenvalues[1] = W1;
  return true;
}

class _InputArray {
 public:
  _InputArray() {}
  _InputArray(const Mat& m) : m_(m) {}
  Mat getMat() const { return m_; }
 private:
  Mat m_;
};
typedef const _InputArray& InputArray;
typedef InputArray InputArrayOfArrays;
class _OutputArray {
 public:
  _OutputArray(Mat& m) : m_(&m) {}
  Mat& getMatRef() const { return *m_; }
 private:
  Mat* m_;
};
typedef const _OutputArray& OutputArray;
inline InputArray noArray() {
  static _InputArray none;
  return none;
}

class Feature2D {
 public:
  virtual ~Feature2D() {}
  virtual void detect(InputArray, std::vector<KeyPoint>&, InputArray = noArray()) {}
  virtual void compute(InputArray, std::vector<KeyPoint>&, OutputArray) {}
  virtual void detectAndCompute(InputArray, InputArray, std::vector<KeyPoint>&, OutputArray, bool = false) {}
  virtual int descriptorSize() const { return 0; }
  virtual int descriptorType() const { return 0; }
};
// Declarations only: brisk.h pulls in headers of classes that okvis never
// constructs (BruteForceMatcher, BriskFeature); they must parse, not run.
struct DMatch {
  int queryIdx, trainIdx, imgIdx;
  float distance;
};
template <typename T> using Ptr = std::shared_ptr<T>;
class DescriptorMatcher {
 public:
  virtual ~DescriptorMatcher() {}
  virtual bool isMaskSupported() const = 0;
  virtual Ptr<DescriptorMatcher> clone(bool emptyTrainData = false) const = 0;
 protected:
  virtual void knnMatchImpl(InputArray, std::vector<std::vector<DMatch> >&, int, InputArrayOfArrays = noArray(),
                            bool = false) = 0;
  virtual void radiusMatchImpl(InputArray, std::vector<std::vector<DMatch> >&, float,
                               InputArrayOfArrays = noArray(), bool = false) = 0;
};
typedef Feature2D FeatureDetector;
typedef Feature2D DescriptorExtractor;

inline void imshow(const std::string&, const Mat&) {}
inline int waitKey(int = 0) { return -1; }

}  // namespace cv

#endif  // ORACLE_CVSHIM_FEATURES2D_HPP_
