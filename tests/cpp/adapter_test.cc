// TEST PROGRAM (not product): compiles adapter/okvis_cuda_adapter.hpp against the REFERENCE's own brisk headers and
// runs it next to the reference's classes, through the same base-class pointers okvis::Frame holds
// (okvis_cv/include/okvis/Frame.hpp:231-232; implementation/Frame.hpp:132, 147):
//   std::shared_ptr<cv::FeatureDetector>     brisk::ScaleSpaceFeatureDetector<HarrisScoreCalculator> | okvis::cuda::BriskDetector
//   std::shared_ptr<cv::DescriptorExtractor> brisk::BriskDescriptorExtractor                         | okvis::cuda::BriskExtractor
// with the camera-awareness maps and the extraction direction handed over through static_pointer_cast to
// brisk::BriskDescriptorExtractor, exactly as Frontend::detectAndDescribe does (Frontend.cpp:214-238).
// Links oracle/_ref/libbrisk_ref.so (the reference's external/brisk sources compiled unchanged) and libokvis_cuda.so.
// Built by __graft_entry__.build() when the reference tree is present; run by tests/test_gpu_adapter.py.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#include <okvis_cuda_adapter.hpp>

#include <brisk/harris-score-calculator.h>
#include <brisk/scale-space-feature-detector.h>

namespace {

struct Rng {   // xorshift64*: the images only have to be the same for both implementations
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 1) {}
  uint32_t next() {
    s ^= s >> 12;
    s ^= s << 25;
    s ^= s >> 27;
    return uint32_t((s * 0x2545F4914F6CDD1Dull) >> 32);
  }
  int below(int n) { return int(next() % uint32_t(n)); }
};

// grey rectangles + a little noise: corners for the Harris detector
cv::Mat scene(int W, int H, uint64_t seed, int rects) {
  cv::Mat img(H, W, CV_8UC1);
  std::memset(img.data, 128, size_t(W) * H);
  Rng rng(seed);
  for (int r = 0; r < rects; ++r) {
    const int x0 = rng.below(W), y0 = rng.below(H), w = 6 + rng.below(90), h = 6 + rng.below(90), v = rng.below(256);
    for (int y = y0; y < std::min(H, y0 + h); ++y) std::memset(img.data + size_t(y) * W + x0, v, size_t(std::min(W, x0 + w) - x0));
  }
  for (size_t i = 0; i < size_t(W) * H; ++i) {
    const int v = int(img.data[i]) + rng.below(7) - 3;
    img.data[i] = uint8_t(v < 0 ? 0 : v > 255 ? 255 : v);
  }
  return img;
}

// maps of an undistorted pinhole camera, as PinholeCamera::getCameraAwarenessMaps lays them out (CV_32FC3 unit rays,
// CV_32FC6 = row-major 2x3 Jacobian of the projection at the ray)
void pinholeMaps(int W, int H, double f, double cu, double cv_, cv::Mat* rays, cv::Mat* jac) {
  rays->create(H, W, CV_32FC3);
  jac->create(H, W, CV_32FC6);
  for (int v = 0; v < H; ++v)
    for (int u = 0; u < W; ++u) {
      const double x = (u - cu) / f, y = (v - cv_) / f, n = std::sqrt(x * x + y * y + 1.0);
      const double X = x / n, Y = y / n, Z = 1.0 / n;
      float* r = reinterpret_cast<float*>(rays->data) + (size_t(v) * W + u) * 3;
      float* j = reinterpret_cast<float*>(jac->data) + (size_t(v) * W + u) * 6;
      r[0] = float(X); r[1] = float(Y); r[2] = float(Z);
      j[0] = float(f / Z); j[1] = 0.f; j[2] = float(-f * X / (Z * Z));
      j[3] = 0.f; j[4] = float(f / Z); j[5] = float(-f * Y / (Z * Z));
    }
}

struct Pair {
  std::shared_ptr<cv::FeatureDetector> detector;
  std::shared_ptr<cv::DescriptorExtractor> extractor;
};

// what okvis::Frame::detect + describe do with the two pointers
void frame(const Pair& p, const cv::Mat& image, const cv::Vec3f& direction, std::vector<cv::KeyPoint>* kps, cv::Mat* desc) {
  std::static_pointer_cast<brisk::BriskDescriptorExtractor>(p.extractor)->setExtractionDirection(direction);
  kps->clear();
  p.detector->detect(image, *kps);
  p.extractor->compute(image, *kps, *desc);
}

int compare(const char* what, const std::vector<cv::KeyPoint>& a, const cv::Mat& da, const std::vector<cv::KeyPoint>& b,
            const cv::Mat& db) {
  if (a.size() != b.size()) {
    std::printf("FAIL %s: %zu reference keypoints, %zu adapter keypoints\n", what, a.size(), b.size());
    return 1;
  }
  double worst = 0.0;
  for (size_t i = 0; i < a.size(); ++i) {
    if (a[i].response != b[i].response || a[i].octave != b[i].octave || a[i].class_id != b[i].class_id ||
        a[i].size != b[i].size) {
      std::printf("FAIL %s: keypoint %zu differs in response / octave / class_id / size\n", what, i);
      return 1;
    }
    worst = std::max(worst, double(std::fabs(a[i].pt.x - b[i].pt.x)));
    worst = std::max(worst, double(std::fabs(a[i].pt.y - b[i].pt.y)));
    worst = std::max(worst, double(std::fabs(a[i].angle - b[i].angle)));
  }
  if (worst > 1e-4) {   // BASELINE.json: floats within 1e-4
    std::printf("FAIL %s: keypoint floats differ by %g\n", what, worst);
    return 1;
  }
  if (!a.empty() && (da.rows != db.rows || da.cols != db.cols || std::memcmp(da.data, db.data, a.size() * 48) != 0)) {
    std::printf("FAIL %s: descriptors differ\n", what);
    return 1;
  }
  std::printf("ok   %s: %zu keypoints, descriptors bit-equal, max float difference %g\n", what, a.size(), worst);
  return 0;
}

template <typename F>
bool throwsRuntimeError(F f) {
  try {
    f();
  } catch (const std::runtime_error&) {
    return true;
  } catch (...) {
    return false;
  }
  return false;
}

}  // namespace

int main() {
  int failures = 0;
  auto device = std::make_shared<okvis::cuda::Device>(0);
  struct Config { const char* name; int W, H; double radius, threshold; size_t maxKp; int rects; };
  const Config configs[] = {{"euroc 752x480 r38 t150 n700", 752, 480, 38.0, 150.0, 700, 500},
                            {"hilti 720x540 r50 t20 n700", 720, 540, 50.0, 20.0, 700, 2500},
                            {"d455 848x480 r30 t5 n2500", 848, 480, 30.0, 5.0, 2500, 5000}};
  for (const Config& c : configs) {
    // Frontend::initialiseBriskFeatureDetectors (Frontend.cpp:2637-2656), reference and replacement
    Pair ref{std::shared_ptr<cv::FeatureDetector>(new brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator>(
                 c.radius, 0, c.threshold, c.maxKp)),
             std::shared_ptr<cv::DescriptorExtractor>(new brisk::BriskDescriptorExtractor(true, false))};
    auto camera = std::make_shared<okvis::cuda::Camera>(device, c.maxKp);
    Pair gpu{std::shared_ptr<cv::FeatureDetector>(new okvis::cuda::BriskDetector(camera, c.radius, 0, c.threshold, c.maxKp)),
             std::shared_ptr<cv::DescriptorExtractor>(new okvis::cuda::BriskExtractor(camera, true, false))};
    std::vector<cv::KeyPoint> ka, kb;
    cv::Mat da, db;
    char what[160];
    // 1. before the maps are set: classic extraction
    cv::Mat img0 = scene(c.W, c.H, 11, c.rects);
    frame(ref, img0, cv::Vec3f(0.f, 1.f, 0.f), &ka, &da);
    frame(gpu, img0, cv::Vec3f(0.f, 1.f, 0.f), &kb, &db);
    std::snprintf(what, sizeof what, "%s, no maps", c.name);
    failures += compare(what, ka, da, kb, db);
    // 2. camera-aware extraction, two frames with different gravity directions
    cv::Mat rays, jac;
    pinholeMaps(c.W, c.H, 455.0, c.W * 0.49, c.H * 0.52, &rays, &jac);
    for (const Pair* p : {&ref, &gpu}) {
      auto x = std::static_pointer_cast<brisk::BriskDescriptorExtractor>(p->extractor);
      if (!x->isCameraAware()) x->setCameraProperties(rays, jac, 455.0f);   // Frontend.cpp:214-231
    }
    for (int f = 0; f < 2; ++f) {
      cv::Mat img = scene(c.W, c.H, 20 + f, c.rects);
      const cv::Vec3f dir = f == 0 ? cv::Vec3f(0.05f, 0.99f, 0.1f) : cv::Vec3f(-0.3f, 0.9f, 0.2f);
      frame(ref, img, dir, &ka, &da);
      frame(gpu, img, dir, &kb, &db);
      std::snprintf(what, sizeof what, "%s, camera-aware frame %d", c.name, f);
      failures += compare(what, ka, da, kb, db);
    }
  }
  // error behaviour of the detector interface (scale-space-feature-detector.h:95-122)
  {
    auto camera = std::make_shared<okvis::cuda::Camera>(device, 700);
    okvis::cuda::BriskDetector det(camera, 38.0, 0, 150.0, 700);
    brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator> refDet(38.0, 0, 150.0, 700);
    std::vector<cv::KeyPoint> k;
    cv::Mat colour(48, 64, CV_32FC3), grey = scene(64, 48, 3, 10), mask = cv::Mat::zeros(48, 64, CV_8UC1), badMask(5, 5, CV_8UC1), none;
    const bool a = throwsRuntimeError([&] { det.detect(colour, k); }) && throwsRuntimeError([&] { refDet.detect(colour, k); });
    const bool b = throwsRuntimeError([&] { det.detect(grey, k, mask); }) && throwsRuntimeError([&] { refDet.detect(grey, k, mask); });
    const bool c = throwsRuntimeError([&] { det.detect(grey, k, badMask); }) && throwsRuntimeError([&] { refDet.detect(grey, k, badMask); });
    det.detect(none, k);   // empty image: returns without touching the keypoints, like the reference (:98-100)
    const bool d = k.empty();
    std::printf("%s errors: wrong image type %d, mask %d, mask of the wrong size %d, empty image %d\n",
                a && b && c && d ? "ok  " : "FAIL", a, b, c, d);
    failures += !(a && b && c && d);
  }
  std::printf(failures ? "ADAPTER_TEST FAIL (%d)\n" : "ADAPTER_TEST PASS\n", failures);
  return failures ? 1 : 0;
}
