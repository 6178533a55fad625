// TEST INFRASTRUCTURE ONLY (oracle). C entry points around the REFERENCE's own
// external/brisk sources, which are compiled unchanged from /root/reference by
// oracle/Makefile into oracle/_ref/libbrisk_ref.so (git-ignored). This file
// contains no algorithm: it only instantiates the reference classes
//   brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator>
//     (external/brisk/include/brisk/scale-space-feature-detector.h:69-170)
//   brisk::BriskDescriptorExtractor
//     (external/brisk/include/brisk/brisk-descriptor-extractor.h:63-200)
//   brisk::Hamming (external/brisk/include/brisk/internal/hamming.h:59-110)
// the way okvis::Frontend does (okvis_frontend/src/Frontend.cpp:2637-2656) and
// copies their results into plain buffers. Used by tests/ and bench.py's
// cpu_baseline / --impl reference legs only.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

#include <brisk/brisk-descriptor-extractor.h>
#include <brisk/harris-score-calculator.h>
#include <brisk/internal/hamming.h>
#include <brisk/internal/harris-scores.h>
#include <brisk/scale-space-feature-detector.h>

namespace brisk {
// Defined (non-inline) in internal/integral-image.h, which brisk-descriptor-extractor.cc already includes.
void IntegralImage8(const cv::Mat& src, cv::Mat* dest);
}  // namespace brisk

namespace {
typedef brisk::HarrisScoreCalculator::PointWithScore Pws;
static_assert(sizeof(Pws) == 8, "PointWithScore layout {int score; u16 x, y}");

struct LayerPeek : public brisk::ScaleSpaceLayer<brisk::HarrisScoreCalculator> {
  const cv::Mat& lut() const { return _LUT; }
};

struct ExtractorPeek : public brisk::BriskDescriptorExtractor {
  ExtractorPeek(bool r, bool s) : brisk::BriskDescriptorExtractor(r, s) {}
  const brisk::BriskPatternPoint* pattern() const { return patternPoints_; }
  unsigned points() const { return points_; }
  const float* scaleList() const { return scaleList_; }
  const unsigned* sizeList() const { return sizeList_; }
  const brisk::BriskShortPair* shortPairs() const { return shortPairs_; }
  const brisk::BriskLongPair* longPairs() const { return longPairs_; }
  unsigned nShort() const { return noShortPairs_; }
  unsigned nLong() const { return noLongPairs_; }
};

cv::Mat wrap_u8(const uint8_t* img, int rows, int cols) {
  return cv::Mat(rows, cols, CV_8UC1, const_cast<uint8_t*>(img));
}
}  // namespace

extern "C" {

int ref_harris(const uint8_t* img, int rows, int cols, int32_t* scores) {
  cv::Mat src = wrap_u8(img, rows, cols), out;
  brisk::HarrisScoresSSE(src, out);
  std::memcpy(scores, out.data, sizeof(int32_t) * static_cast<size_t>(rows) * cols);
  return 0;
}

// Raster-order candidate list (HarrisScoreCalculator::Get2dMaxima).
int ref_maxima(const uint8_t* img, int rows, int cols, int absThreshold, void* out, int cap) {
  brisk::HarrisScoreCalculator calc;
  calc.SetImage(wrap_u8(img, rows, cols), true);
  std::vector<Pws> pts;
  calc.Get2dMaxima(pts, absThreshold);
  int n = static_cast<int>(pts.size());
  std::memcpy(out, pts.data(), sizeof(Pws) * static_cast<size_t>(n < cap ? n : cap));
  return n;
}

// std::sort + greedy uniformity on a candidate list, layer-0 LUT.
int ref_uniformity(void* inout, int n, double radius, int rows, int cols, uint64_t maxNumKpt) {
  std::vector<uint8_t> tiny(32 * 32, 0);
  LayerPeek layer;
  layer.Create(wrap_u8(tiny.data(), 32, 32), false);
  Pws* p = static_cast<Pws*>(inout);
  std::vector<Pws> pts(p, p + n);
  if (pts.empty()) return 0;
  EnforceKeyPointUniformity(layer.lut(), radius, rows, cols, static_cast<size_t>(maxNumKpt), pts);
  std::memcpy(inout, pts.data(), sizeof(Pws) * pts.size());
  return static_cast<int>(pts.size());
}

int ref_lut(float* out) {
  std::vector<uint8_t> tiny(32 * 32, 0);
  LayerPeek layer;
  layer.Create(wrap_u8(tiny.data(), 32, 32), false);
  std::memcpy(out, layer.lut().data, sizeof(float) * 31 * 31);
  return 0;
}

// Full detector as okvis constructs it (Frontend.cpp:2645-2648).
// Returns the number of keypoints, or -1 when the reference threw.
int ref_detect(const uint8_t* img, int rows, int cols, double radius, uint64_t octaves,
               double absThreshold, uint64_t maxNumKpt, void* keypoints, int cap) {
  try {
    brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator> det(
        radius, static_cast<size_t>(octaves), absThreshold, static_cast<size_t>(maxNumKpt));
    std::vector<cv::KeyPoint> kps;
    det.detect(wrap_u8(img, rows, cols), kps);
    int n = static_cast<int>(kps.size());
    std::memcpy(keypoints, kps.data(), sizeof(cv::KeyPoint) * static_cast<size_t>(n < cap ? n : cap));
    return n;
  } catch (const std::exception&) {
    return -1;
  }
}

void* ref_extractor_create(int rotationInvariant, int scaleInvariant) {
  return new ExtractorPeek(rotationInvariant != 0, scaleInvariant != 0);
}
void ref_extractor_destroy(void* h) { delete static_cast<ExtractorPeek*>(h); }

// Copies the maps (the reference keeps cv::Mat references to caller memory).
void ref_extractor_set_camera(void* h, const float* rays, const float* jacobians, int rows, int cols,
                              float virtualFocalLength) {
  cv::Mat r(rows, cols, CV_32FC3), j(rows, cols, CV_32FC6);
  std::memcpy(r.data, rays, sizeof(float) * 3 * static_cast<size_t>(rows) * cols);
  std::memcpy(j.data, jacobians, sizeof(float) * 6 * static_cast<size_t>(rows) * cols);
  static_cast<ExtractorPeek*>(h)->setCameraProperties(r, j, virtualFocalLength);
}
void ref_extractor_set_direction(void* h, const float* dir) {
  static_cast<ExtractorPeek*>(h)->setExtractionDirection(cv::Vec3f(dir[0], dir[1], dir[2]));
}

// keypoints: in/out (border filter shrinks it; angle is written back).
int ref_describe(void* h, const uint8_t* img, int rows, int cols, void* keypoints, int n, uint8_t* desc) {
  try {
    cv::KeyPoint* kp = static_cast<cv::KeyPoint*>(keypoints);
    std::vector<cv::KeyPoint> kps(kp, kp + n);
    cv::Mat d;
    static_cast<ExtractorPeek*>(h)->compute(wrap_u8(img, rows, cols), kps, d);
    std::memcpy(kp, kps.data(), sizeof(cv::KeyPoint) * kps.size());
    if (!kps.empty()) std::memcpy(desc, d.data, 48 * kps.size());
    return static_cast<int>(kps.size());
  } catch (const std::exception&) {
    return -1;
  }
}

int ref_integral(const uint8_t* img, int rows, int cols, int32_t* out) {
  cv::Mat integral;
  brisk::IntegralImage8(wrap_u8(img, rows, cols), &integral);
  std::memcpy(out, integral.data, sizeof(int32_t) * static_cast<size_t>(rows + 1) * (cols + 1));
  return 0;
}

// Pattern tables built by the reference constructor (for pinning our generator).
int ref_pattern_points(void* h, int scale, float* out /* 1024*66*3 */) {
  ExtractorPeek* e = static_cast<ExtractorPeek*>(h);
  size_t per = static_cast<size_t>(1024) * e->points();
  std::memcpy(out, e->pattern() + per * scale, per * sizeof(brisk::BriskPatternPoint));
  return static_cast<int>(e->points());
}
int ref_pattern_meta(void* h, float* scaleList64, uint32_t* sizeList64, uint32_t* shortPairs /*384*2*/,
                     int32_t* longPairs /* n*4 */, int longCap) {
  ExtractorPeek* e = static_cast<ExtractorPeek*>(h);
  std::memcpy(scaleList64, e->scaleList(), 64 * sizeof(float));
  std::memcpy(sizeList64, e->sizeList(), 64 * sizeof(uint32_t));
  for (unsigned p = 0; p < e->nShort(); ++p) {
    shortPairs[2 * p] = e->shortPairs()[p].i;
    shortPairs[2 * p + 1] = e->shortPairs()[p].j;
  }
  int n = static_cast<int>(e->nLong());
  for (int p = 0; p < n && p < longCap; ++p) {
    longPairs[4 * p] = static_cast<int32_t>(e->longPairs()[p].i);
    longPairs[4 * p + 1] = static_cast<int32_t>(e->longPairs()[p].j);
    longPairs[4 * p + 2] = e->longPairs()[p].weighted_dx;
    longPairs[4 * p + 3] = e->longPairs()[p].weighted_dy;
  }
  return n;
}

uint32_t ref_hamming(const uint8_t* a, const uint8_t* b, int numberOf128BitWords) {
  return brisk::Hamming::PopcntofXORed(a, b, numberOf128BitWords);
}

// Best match per query, the way every matcher loop in Frontend.cpp uses the
// primitive: strict '<' running minimum, first candidate wins ties.
// best[q] = (dist << 32) | index, or ~0 when nothing is below the threshold.
void ref_hamming_best(const uint8_t* queries, int nq, const uint8_t* map, int nm, uint32_t threshold,
                      uint64_t* best) {
  for (int q = 0; q < nq; ++q) {
    uint32_t bd = threshold;
    uint32_t bi = 0xffffffffu;
    const uint8_t* qd = queries + 48 * static_cast<size_t>(q);
    for (int m = 0; m < nm; ++m) {
      uint32_t d = brisk::Hamming::PopcntofXORed(qd, map + 48 * static_cast<size_t>(m), 3);
      if (d < bd) {
        bd = d;
        bi = static_cast<uint32_t>(m);
      }
    }
    best[q] = bi == 0xffffffffu ? ~0ull : (static_cast<uint64_t>(bd) << 32) | bi;
  }
}

// The three OpenCV-defined pieces of arithmetic on the extractor path live in the cv shim (OpenCV's C++ sources
// are neither in /root/reference nor in this image). These entry points expose the shim's versions so that
// tests/test_oracle_pins.py can pin them against goldens produced by the Python cv2 4.13 wheel
// (tests/golden/make_golden.py): cv::eigen (brisk-descriptor-extractor.cc:716-719), the 2x3 * 3x2 Mat product
// (:712) and cv::normalize(Vec3f) (:689-690).
void shim_eigen2x2(const float* a /*4, row-major*/, float* ev /*2*/) {
  cv::Mat m(2, 2, CV_32FC1);
  std::memcpy(m.data, a, 4 * sizeof(float));
  cv::Vec2f e;
  cv::eigen(m, e);
  ev[0] = e[0];
  ev[1] = e[1];
}
void shim_gemm_2x3_3x2(const float* a /*6*/, const float* b /*6*/, float* out /*4*/) {
  cv::Mat A(2, 3, CV_32FC1), B(3, 2, CV_32FC1);
  std::memcpy(A.data, a, 6 * sizeof(float));
  std::memcpy(B.data, b, 6 * sizeof(float));
  cv::Mat d = A * B;
  std::memcpy(out, d.data, 4 * sizeof(float));
}
void shim_normalize3(const float* v, float* out) {
  cv::Vec3f r = cv::normalize(cv::Vec3f(v[0], v[1], v[2]));
  out[0] = r[0];
  out[1] = r[1];
  out[2] = r[2];
}

// std::sort exactly as EnforceKeyPointUniformity calls it (uniformity-enforcement-inl.h:67): the reference's own
// PointWithScore::operator< (score-calculator.h:94-96) under this toolchain's libstdc++ introsort. The order of
// equal scores is what the device emulation has to reproduce.
void ref_std_sort(void* inout, int n) {
  Pws* p = static_cast<Pws*>(inout);
  std::sort(p, p + n);
}

}  // extern "C"
