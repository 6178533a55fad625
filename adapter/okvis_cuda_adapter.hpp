// OKVIS2-X side of the okvis_cuda drop-in: the two OpenCV virtual interfaces okvis::Frame calls
// (okvis_cv/include/okvis/implementation/Frame.hpp:132 detector_->detect, :147 extractor_->compute), implemented on
// libokvis_cuda.so (include/okvis_cuda.h). A maintainer copies this header into okvis_frontend/include/okvis/ and
// changes the two `new` expressions of Frontend::initialiseBriskFeatureDetectors (okvis_frontend/src/Frontend.cpp:
// 2645-2652); nothing else in the reference has to know.
//
// Needs only <opencv2/features2d/features2d.hpp>, the reference's <brisk/brisk-descriptor-extractor.h> and
// okvis_cuda.h. It is compiled and run in this repository against the reference's own brisk headers and the compiled
// reference classes (tests/cpp/adapter_test.cc, built by __graft_entry__.build(), run by pytest -m gpu).
//
//   brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator>  ->  okvis::cuda::BriskDetector
//       (external/brisk/include/brisk/scale-space-feature-detector.h:69-170)
//   brisk::BriskDescriptorExtractor                                  ->  okvis::cuda::BriskExtractor
//       (external/brisk/include/brisk/brisk-descriptor-extractor.h:63-200); derived from it because
//       Frontend::detectAndDescribe static-casts the extractor to that type to hand over the camera-awareness maps
//       and the extraction direction (Frontend.cpp:214-238)
#ifndef OKVIS_CUDA_ADAPTER_HPP_
#define OKVIS_CUDA_ADAPTER_HPP_

#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/features2d/features2d.hpp>
#include <brisk/brisk-descriptor-extractor.h>

#include <okvis_cuda.h>

namespace okvis {
namespace cuda {

static_assert(sizeof(cv::KeyPoint) == sizeof(okc_keypoint), "cv::KeyPoint and okc_keypoint share one 28-byte layout");

/// The reference throws std::runtime_error for bad arguments (scale-space-feature-detector.h:95-122); so does this.
inline void check(int rc, const char* what) {
  if (rc == OKC_OK) return;
  throw std::runtime_error(std::string(what) + ": " + okc_last_error_string());
}

/// One per process and device; shared by the cameras.
class Device {
 public:
  explicit Device(int device = -1) { check(okc_context_create(device, &ctx_), "okc_context_create"); }
  ~Device() { okc_context_destroy(ctx_); }
  Device(const Device&) = delete;
  Device& operator=(const Device&) = delete;
  okc_context* get() const { return ctx_; }

 private:
  okc_context* ctx_ = nullptr;
};

/// Device side of one camera index: stream, scratch, camera-awareness maps. Shared by the detector and the extractor
/// of that camera (the reference keeps one detector / extractor pair per camera behind featureDetectorMutexes_,
/// Frontend.cpp:208, 2637-2656), created lazily because image size and maps are only known at the first frame.
class Camera {
 public:
  Camera(std::shared_ptr<Device> device, size_t maxNumKeypoints)
      : device_(std::move(device)), capacity_(int(maxNumKeypoints) + 64) {}
  ~Camera() {
    if (cam_) okc_camera_destroy(cam_);
  }
  Camera(const Camera&) = delete;
  Camera& operator=(const Camera&) = delete;

  /// The detector's view: any device camera of the image's shape will do (detection does not read the maps).
  okc_camera* ensure(const cv::Mat& image) {
    if (cam_ && width_ == image.cols && height_ == image.rows) return cam_;
    return ensure(image, nullptr, nullptr, 500.0f);
  }

  /// The extractor's view: (re)creates the device camera when the image shape or the maps change (the maps are
  /// uploaded once, at creation). rays: CV_32FC3, jacobians: CV_32FC6, as getCameraAwarenessMaps() returns them.
  okc_camera* ensure(const cv::Mat& image, const cv::Mat* rays, const cv::Mat* jacobians, float virtualFocalLength) {
    const void* raysKey = rays ? static_cast<const void*>(rays->data) : nullptr;
    if (cam_ && width_ == image.cols && height_ == image.rows && raysKey == raysKey_ &&
        (rays == nullptr || focal_ == virtualFocalLength))
      return cam_;
    if (rays && (rays->rows != image.rows || rays->cols != image.cols || jacobians == nullptr ||
                 jacobians->rows != image.rows || jacobians->cols != image.cols))
      throw std::runtime_error("camera-awareness maps must have the image dimensions");
    if (cam_) {
      okc_camera_destroy(cam_);
      cam_ = nullptr;
    }
    check(okc_camera_create(device_->get(), image.cols, image.rows,
                            rays ? reinterpret_cast<const float*>(rays->data) : nullptr,
                            rays ? reinterpret_cast<const float*>(jacobians->data) : nullptr, virtualFocalLength,
                            capacity_, &cam_),
          "okc_camera_create");
    width_ = image.cols;
    height_ = image.rows;
    raysKey_ = raysKey;
    focal_ = virtualFocalLength;
    return cam_;
  }
  int capacity() const { return capacity_; }
  std::mutex& mutex() { return mutex_; }

 private:
  std::shared_ptr<Device> device_;
  okc_camera* cam_ = nullptr;
  int capacity_;
  int width_ = 0, height_ = 0;
  const void* raysKey_ = nullptr;
  float focal_ = 0.f;
  std::mutex mutex_;
};

/// Drop-in for brisk::ScaleSpaceFeatureDetector<brisk::HarrisScoreCalculator>: same constructor arguments, same
/// exceptions, same keypoints in the same order (tests/test_gpu_detect_describe.py, tests/cpp/adapter_test.cc).
class BriskDetector : public cv::FeatureDetector {
 public:
  BriskDetector(std::shared_ptr<Camera> camera, double uniformityRadius, size_t octaves, double absoluteThreshold = 0,
                size_t maxNumKpt = 2000)
      : camera_(std::move(camera)) {
    params_.uniformity_radius = uniformityRadius;
    params_.octaves = octaves;
    params_.absolute_threshold = absoluteThreshold;
    params_.max_num_keypoints = maxNumKpt;
  }

  void detect(cv::InputArray image_, std::vector<cv::KeyPoint>& keypoints,
              cv::InputArray mask_ = cv::noArray()) override {
    cv::Mat image = image_.getMat();
    cv::Mat mask = mask_.getMat();
    if (!image.isContinuous() || image.channels() != 1 || image.elemSize() != 1)
      throw std::runtime_error("BRISK requires continuous 1-channel 8-bit images");   // :95-97
    if (image.empty()) return;                                                          // :98-100
    if (!(mask.empty() || (mask.type() == CV_8UC1 && mask.rows == image.rows && mask.cols == image.cols)))
      throw std::runtime_error("mask must be empty or CV_8UC1 type with image dimensions");   // :101-104
    if (!mask.empty()) throw std::runtime_error("BRISK does not support masks yet");    // :119-122
    // the reference re-scores passed keypoints instead of detecting (:124-128); okvis never does that
    // (Frame::detect clears them, Frame.hpp:123; Frontend.cpp:212 refuses external keypoints)
    if (!keypoints.empty()) throw std::runtime_error("okvis_cuda: passed keypoints are not supported");
    std::lock_guard<std::mutex> lock(camera_->mutex());
    okc_camera* cam = camera_->ensure(image);
    keypoints.resize(size_t(camera_->capacity()));
    int n = 0;
    const int rc = okc_detect(cam, image.data, &params_, reinterpret_cast<okc_keypoint*>(keypoints.data()),
                              camera_->capacity(), &n);
    keypoints.resize(rc == OKC_OK ? size_t(n) : 0);
    check(rc, "okc_detect");
  }

 private:
  std::shared_ptr<Camera> camera_;
  okc_detect_params params_;
};

/// Drop-in for brisk::BriskDescriptorExtractor as okvis constructs it (rotation invariant, not scale invariant,
/// Frontend.cpp:2650-2652). setCameraProperties / setExtractionDirection / isCameraAware are the inherited,
/// non-virtual members (brisk-descriptor-extractor.h:112-122): the maps and the direction they store are read here.
class BriskExtractor : public brisk::BriskDescriptorExtractor {
 public:
  BriskExtractor(std::shared_ptr<Camera> camera, bool rotationInvariant = true, bool scaleInvariant = false)
      : brisk::BriskDescriptorExtractor(rotationInvariant, scaleInvariant), camera_(std::move(camera)) {
    if (scaleInvariant) throw std::runtime_error("okvis_cuda: scale-invariant extraction is not supported");
  }

  void compute(cv::InputArray image_, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors) override {
    cv::Mat image = image_.getMat();
    if (!image.isContinuous() || image.channels() != 1 || image.elemSize() != 1)
      throw std::runtime_error("BRISK requires continuous 1-channel 8-bit images");
    cv::Mat out;
    std::lock_guard<std::mutex> lock(camera_->mutex());
    const bool aware = isCameraAware();
    okc_camera* cam = camera_->ensure(image, aware ? &rayDirections_ : nullptr, aware ? &imageJacobians_ : nullptr,
                                      virtualFocalLength_);
    okc_describe_params p;
    p.rotation_invariant = rotationInvariance ? 1 : 0;
    p.scale_invariant = scaleInvariance ? 1 : 0;
    for (int i = 0; i < 3; ++i) p.extraction_direction[i] = extractionDirection_[i];
    if (int(keypoints.size()) > camera_->capacity())
      throw std::runtime_error("okvis_cuda: more keypoints than the camera was created for");
    int n = int(keypoints.size());
    out.create(n > 0 ? n : 1, OKC_DESC_BYTES, CV_8UC1);
    const int rc = okc_describe(cam, image.data, &p, reinterpret_cast<okc_keypoint*>(keypoints.data()), &n, out.data);
    check(rc, "okc_describe");
    keypoints.resize(size_t(n));                    // keypoints too close to the border are removed, in order
    cv::Mat result;
    if (n > 0) {
      result.create(n, OKC_DESC_BYTES, CV_8UC1);
      std::memcpy(result.data, out.data, size_t(n) * OKC_DESC_BYTES);
    }
    descriptors.getMatRef() = result;
  }

 private:
  std::shared_ptr<Camera> camera_;
};

}  // namespace cuda
}  // namespace okvis

#endif  // OKVIS_CUDA_ADAPTER_HPP_
