// C ABI of okvis_cuda (include/okvis_cuda.h): handles, staging, launches. No algorithm lives here.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"
#include "pattern.h"

namespace okc {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// growable device / pinned-host scratch
struct Buf {
  void* p = nullptr;
  size_t cap = 0;
  bool pinned = false;
  int ensure(size_t bytes) {
    if (bytes <= cap) return OKC_OK;
    release();
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      p = nullptr;
      cap = 0;
      set_error("allocation of %zu bytes failed: %s", want, cudaGetErrorString(e));
      return OKC_ERR_CUDA;
    }
    cap = want;
    return OKC_OK;
  }
  void release() {
    if (p) {
      if (pinned) cudaFreeHost(p);
      else cudaFree(p);
    }
    p = nullptr;
    cap = 0;
  }
};

}  // namespace okc

using namespace okc;

struct okc_context {
  int device = 0;
  cudaStream_t stream = nullptr;
  PatternTables tables{};
  HostTables host;
  std::mutex mu;        // matching entry points share the scratch below
  Buf d[8];             // device scratch
  Buf h[4];             // pinned host scratch
  okc_context() {
    for (auto& b : h) b.pinned = true;
  }
};

struct okc_camera {
  okc_context* ctx = nullptr;
  int W = 0, H = 0, max_kp = 0;
  bool camera_aware = false;
  float focal = 500.0f;
  cudaStream_t stream = nullptr;
  float* rays = nullptr;
  float* jac = nullptr;
  DeviceImageScratch s{};
  // pinned staging
  uint8_t* h_image = nullptr;
  okc_keypoint* h_kp = nullptr;
  uint8_t* h_desc = nullptr;
  int* h_counters = nullptr;
  bool pending = false;
  // frame pipeline (capi_frame.inc): map-match outputs of the frame in flight
  double* d_proj = nullptr;       // 2 * proj_cap doubles
  double* h_proj = nullptr;       // pinned copy (the event replay needs the projections on the host)
  int proj_cap = 0;
  uint8_t* d_match = nullptr;     // [best_dist | best_owner | events | count] block
  uint8_t* h_match = nullptr;
  size_t match_bytes = 0;
  int event_cap = 0;
  int match_landmarks = -1;       // >= 0: a map match is in flight for that many landmarks
  bool proj_dirty_host = false;   // projections were given as a device pointer: fetch them at finish
  // per-kernel timing (okc_debug_kernel_timing)
  static constexpr int kMaxMarks = 10;
  bool timing = false;
  cudaEvent_t ev[kMaxMarks + 1] = {};
  const char* ev_name[kMaxMarks] = {};
  int n_marks = 0;
  void timing_begin() {
    if (!timing) return;
    n_marks = 0;
    cudaEventRecord(ev[0], stream);
  }
  void mark(const char* name) {  // call right after the launch of kernel `name`
    if (!timing || n_marks >= kMaxMarks) return;
    ev_name[n_marks] = name;
    cudaEventRecord(ev[++n_marks], stream);
  }
};

struct okc_descriptor_set {
  okc_context* ctx = nullptr;
  void* rows = nullptr;
  int64_t n = 0;
};

namespace okc {
int launch_hamming_best(const void* queries, int nq, const void* map, long long nm, unsigned threshold,
                        unsigned long long indexOffset, unsigned long long* best, cudaStream_t s);
extern int g_match_prefix;
}  // namespace okc

extern "C" {

const char* okc_last_error_string(void) { return g_err; }
int okc_version(void) { return 100; }

int okc_context_create(int device, okc_context** out) {
  if (!out) return OKC_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    set_error("no CUDA device: %s (okvis_cuda has no CPU fallback)", cudaGetErrorString(e));
    return OKC_ERR_CUDA;
  }
  if (device < 0) OKC_CUDA_TRY(cudaGetDevice(&device));
  OKC_CUDA_TRY(cudaSetDevice(device));
  okc_context* c = new (std::nothrow) okc_context();
  if (!c) return OKC_ERR_CUDA;
  c->device = device;
  OKC_CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  build_host_tables(&c->host);
  const HostTables& t = c->host;
  OKC_CUDA_TRY(cudaMalloc(&c->tables.points, t.points.size() * sizeof(HostPatternPoint)));
  OKC_CUDA_TRY(cudaMalloc(&c->tables.short_pairs, t.short_pairs.size()));
  OKC_CUDA_TRY(cudaMalloc(&c->tables.long_pairs, t.long_pairs.size() * sizeof(HostLongPair)));
  OKC_CUDA_TRY(cudaMalloc(&c->tables.lut, t.lut.size() * sizeof(float)));
  OKC_CUDA_TRY(cudaMemcpy(c->tables.points, t.points.data(), t.points.size() * sizeof(HostPatternPoint),
                          cudaMemcpyHostToDevice));
  OKC_CUDA_TRY(cudaMemcpy(c->tables.short_pairs, t.short_pairs.data(), t.short_pairs.size(), cudaMemcpyHostToDevice));
  OKC_CUDA_TRY(cudaMemcpy(c->tables.long_pairs, t.long_pairs.data(), t.long_pairs.size() * sizeof(HostLongPair),
                          cudaMemcpyHostToDevice));
  OKC_CUDA_TRY(cudaMemcpy(c->tables.lut, t.lut.data(), t.lut.size() * sizeof(float), cudaMemcpyHostToDevice));
  c->tables.border = t.border;
  c->tables.basicscale = t.basicscale;
  *out = c;
  return OKC_OK;
}

int okc_context_destroy(okc_context* c) {
  if (!c) return OKC_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  cudaFree(c->tables.points);
  cudaFree(c->tables.short_pairs);
  cudaFree(c->tables.long_pairs);
  cudaFree(c->tables.lut);
  for (auto& b : c->d) b.release();
  for (auto& b : c->h) b.release();
  cudaStreamDestroy(c->stream);
  delete c;
  return OKC_OK;
}

int okc_context_device(const okc_context* c) { return c ? c->device : -1; }

// Host tables only (no device work): lets CPU-only tests pin pattern.cc against the reference constructor.
int okc_debug_host_tables(float* points /*1024*66*3*/, int* basicscale, int* border, int32_t* long_pairs /*856*4*/,
                          uint8_t* short_pairs /*384*2*/, float* lut /*31*31*/) {
  HostTables t;
  build_host_tables(&t);
  if (points) memcpy(points, t.points.data(), t.points.size() * sizeof(HostPatternPoint));
  if (basicscale) *basicscale = t.basicscale;
  if (border) *border = t.border;
  if (long_pairs)
    for (size_t p = 0; p < t.long_pairs.size(); ++p) {
      long_pairs[4 * p] = t.long_pairs[p].i;
      long_pairs[4 * p + 1] = t.long_pairs[p].j;
      long_pairs[4 * p + 2] = t.long_pairs[p].wdx;
      long_pairs[4 * p + 3] = t.long_pairs[p].wdy;
    }
  if (short_pairs) memcpy(short_pairs, t.short_pairs.data(), t.short_pairs.size());
  if (lut) memcpy(lut, t.lut.data(), t.lut.size() * sizeof(float));
  return OKC_OK;
}

int okc_camera_create(okc_context* ctx, int width, int height, const float* rays, const float* jacobians,
                      float virtual_focal_length, int max_keypoints, okc_camera** out) {
  if (!ctx || !out || width < 40 || height < 40 || max_keypoints <= 0 || ((rays == nullptr) != (jacobians == nullptr))) {
    set_error("okc_camera_create: invalid argument");
    return OKC_ERR_INVALID;
  }
  *out = nullptr;
  OKC_CUDA_TRY(cudaSetDevice(ctx->device));
  okc_camera* c = new (std::nothrow) okc_camera();
  if (!c) return OKC_ERR_CUDA;
  c->ctx = ctx;
  c->W = width;
  c->H = height;
  c->max_kp = max_keypoints;
  c->focal = virtual_focal_length;
  const size_t px = (size_t)width * height;
  OKC_CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  OKC_CUDA_TRY(cudaMalloc(&c->s.image, px));
  OKC_CUDA_TRY(cudaMalloc(&c->s.scores, px * sizeof(int32_t)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.cand, ((size_t)kCandCap + max_keypoints) * sizeof(uint64_t)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.bitmap, (size_t)height * ((width + 31) / 32) * sizeof(uint32_t)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.order_scratch, order_scratch_bytes()));
  OKC_CUDA_TRY(cudaMalloc(&c->s.counters, 8 * sizeof(int)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.kp_detect, (size_t)max_keypoints * sizeof(okc_keypoint)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.kp_desc, (size_t)max_keypoints * sizeof(okc_keypoint)));
  OKC_CUDA_TRY(cudaMalloc(&c->s.desc, (size_t)max_keypoints * OKC_DESC_BYTES));
  OKC_CUDA_TRY(cudaMallocHost(&c->h_image, px));
  OKC_CUDA_TRY(cudaMallocHost(&c->h_kp, (size_t)max_keypoints * sizeof(okc_keypoint)));
  OKC_CUDA_TRY(cudaMallocHost(&c->h_desc, (size_t)max_keypoints * OKC_DESC_BYTES));
  OKC_CUDA_TRY(cudaMallocHost(&c->h_counters, 8 * sizeof(int)));
  if (rays) {
    c->camera_aware = true;
    OKC_CUDA_TRY(cudaMalloc(&c->rays, px * 3 * sizeof(float)));
    OKC_CUDA_TRY(cudaMalloc(&c->jac, px * 6 * sizeof(float)));
    OKC_CUDA_TRY(cudaMemcpy(c->rays, rays, px * 3 * sizeof(float), cudaMemcpyHostToDevice));
    OKC_CUDA_TRY(cudaMemcpy(c->jac, jacobians, px * 6 * sizeof(float), cudaMemcpyHostToDevice));
  }
  *out = c;
  return OKC_OK;
}

int okc_camera_destroy(okc_camera* c) {
  if (!c) return OKC_OK;
  cudaSetDevice(c->ctx->device);
  cudaStreamSynchronize(c->stream);
  cudaFree(c->s.image);
  cudaFree(c->s.scores);
  cudaFree(c->s.cand);
  cudaFree(c->s.bitmap);
  cudaFree(c->s.order_scratch);
  cudaFree(c->s.counters);
  cudaFree(c->s.kp_detect);
  cudaFree(c->s.kp_desc);
  cudaFree(c->s.desc);
  cudaFree(c->s.occupancy);
  cudaFree(c->rays);
  cudaFree(c->jac);
  cudaFreeHost(c->h_image);
  cudaFreeHost(c->h_kp);
  cudaFreeHost(c->h_desc);
  cudaFreeHost(c->h_counters);
  cudaFree(c->d_proj);
  cudaFreeHost(c->h_proj);
  cudaFree(c->d_match);
  cudaFreeHost(c->h_match);
  for (auto& e : c->ev)
    if (e) cudaEventDestroy(e);
  cudaStreamDestroy(c->stream);
  delete c;
  return OKC_OK;
}

}  // extern "C"

// ---- shared pieces of detect / describe ----
static int check_detect_params(const okc_camera* cam, const okc_detect_params* p, DetectArgs* a) {
  if (!cam || !p) return OKC_ERR_INVALID;
  if (p->octaves != 0) {
    set_error("octaves > 0 (scale space) is not implemented: every shipped OKVIS2-X config uses 0");
    return OKC_ERR_INVALID;
  }
  double radius = p->uniformity_radius;
  if (radius == 0) radius = 1;  // ScaleSpaceLayer::SetUniformityRadius
  if (!(radius > 0)) {
    set_error("uniformity radius must be >= 0");
    return OKC_ERR_INVALID;
  }
  a->width = cam->W;
  a->height = cam->H;
  a->abs_threshold = (int)p->absolute_threshold;  // double _absoluteThreshold -> int parameter of Get2dMaxima
  a->scaling = (float)(15.0 / (double)(float)radius);
  a->max_kpt = (unsigned long long)(double)p->max_num_keypoints;  // _maxNumKpt / factor(=1.0) -> size_t
  a->max_out = cam->max_kp;
  return OKC_OK;
}

static int ensure_occupancy(okc_camera* cam, const DetectArgs& a) {
  // global fallback for the occupancy image (used when it does not fit in shared memory)
  const size_t rows = (size_t)((float)(a.height - 3) * a.scaling + 16.0f) + 17;
  const size_t stride = (((size_t)((float)(a.width - 3) * a.scaling + 16.0f) + 17) + 15) & ~(size_t)15;
  const size_t bytes = rows * stride + 64;
  if (bytes > cam->s.occupancy_bytes) {
    cudaFree(cam->s.occupancy);
    cam->s.occupancy = nullptr;
    cam->s.occupancy_bytes = 0;
    OKC_CUDA_TRY(cudaMalloc(&cam->s.occupancy, bytes));
    cam->s.occupancy_bytes = bytes;
  }
  return OKC_OK;
}

static int upload_image(okc_camera* cam, const uint8_t* image) {
  const size_t px = (size_t)cam->W * cam->H;
  memcpy(cam->h_image, image, px);
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->s.image, cam->h_image, px, cudaMemcpyHostToDevice, cam->stream));
  return OKC_OK;
}

static int enqueue_detect(okc_camera* cam, const uint8_t* image, const DetectArgs& a) {
  OKC_CUDA_TRY(cudaMemsetAsync(cam->s.counters, 0, 8 * sizeof(int), cam->stream));
  cam->timing_begin();
  launch_harris_nms(image, cam->s.scores, cam->s.bitmap, a, cam->stream);
  cam->mark("harris_nms");
  launch_order_candidates(cam->s.scores, cam->s.bitmap, cam->s.cand, cam->s.counters, cam->s.order_scratch, a,
                          cam->stream);
  cam->mark("order_candidates");
  launch_select_keypoints(cam->s.scores, cam->s.cand, cam->s.counters, cam->s.occupancy, cam->ctx->tables.lut,
                          cam->s.kp_detect, a, cam->stream);
  cam->mark("select_keypoints");
  OKC_CUDA_TRY(cudaGetLastError());
  return OKC_OK;
}

static int enqueue_describe(okc_camera* cam, const uint8_t* image, const okc_describe_params* p,
                            const okc_keypoint* kp_in_dev, const int* n_in_dev) {
  if (p->scale_invariant) {
    set_error("scale_invariant extraction is not implemented: OKVIS2-X constructs the extractor with false");
    return OKC_ERR_INVALID;
  }
  DescribeArgs d;
  d.width = cam->W;
  d.height = cam->H;
  d.rotation_invariant = p->rotation_invariant ? 1 : 0;
  d.camera_aware = cam->camera_aware ? 1 : 0;
  d.dir[0] = p->extraction_direction[0];
  d.dir[1] = p->extraction_direction[1];
  d.dir[2] = p->extraction_direction[2];
  d.inv_f = 1.0f / cam->focal;
  launch_describe(image, kp_in_dev, n_in_dev, cam->s.kp_desc, cam->s.desc, cam->s.counters, cam->rays, cam->jac,
                  cam->ctx->tables, d, cam->max_kp, cam->stream);
  cam->mark("border_filter+describe");
  OKC_CUDA_TRY(cudaGetLastError());
  return OKC_OK;
}

static int check_flags(const okc_camera* cam) {
  if (cam->h_counters[3] == 1) {
    set_error("more than %d NMS candidates in one image", kCandCap);
    return OKC_ERR_CAPACITY;
  }
  if (cam->h_counters[3] == 2) {
    set_error("more keypoints than the camera handle's max_keypoints (%d)", cam->max_kp);
    return OKC_ERR_CAPACITY;
  }
  return OKC_OK;
}

extern "C" {

int okc_detect(okc_camera* cam, const uint8_t* image, const okc_detect_params* p, okc_keypoint* keypoints,
               int capacity, int* n) {
  DetectArgs a;
  int rc = check_detect_params(cam, p, &a);
  if (rc) return rc;
  if (!image || !keypoints || !n) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  if ((rc = ensure_occupancy(cam, a))) return rc;
  if ((rc = upload_image(cam, image))) return rc;
  if ((rc = enqueue_detect(cam, cam->s.image, a))) return rc;
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_counters, cam->s.counters, 8 * sizeof(int), cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_kp, cam->s.kp_detect, (size_t)cam->max_kp * sizeof(okc_keypoint),
                               cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaStreamSynchronize(cam->stream));
  if ((rc = check_flags(cam))) return rc;
  const int cnt = cam->h_counters[1];
  if (cnt > capacity) {
    set_error("okc_detect: %d keypoints exceed the caller's capacity %d", cnt, capacity);
    return OKC_ERR_CAPACITY;
  }
  memcpy(keypoints, cam->h_kp, (size_t)cnt * sizeof(okc_keypoint));
  *n = cnt;
  return OKC_OK;
}

int okc_describe(okc_camera* cam, const uint8_t* image, const okc_describe_params* p, okc_keypoint* keypoints, int* n,
                 uint8_t* descriptors) {
  if (!cam || !image || !p || !keypoints || !n || !descriptors || *n < 0) return OKC_ERR_INVALID;
  if (*n > cam->max_kp) {
    set_error("okc_describe: %d keypoints exceed the camera handle's max_keypoints (%d)", *n, cam->max_kp);
    return OKC_ERR_CAPACITY;
  }
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  int rc;
  if ((rc = upload_image(cam, image))) return rc;
  memcpy(cam->h_kp, keypoints, (size_t)*n * sizeof(okc_keypoint));
  cam->h_counters[4] = *n;
  OKC_CUDA_TRY(cudaMemsetAsync(cam->s.counters, 0, 8 * sizeof(int), cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->s.counters + 1, cam->h_counters + 4, sizeof(int), cudaMemcpyHostToDevice,
                               cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->s.kp_detect, cam->h_kp, (size_t)*n * sizeof(okc_keypoint), cudaMemcpyHostToDevice,
                               cam->stream));
  cam->timing_begin();
  if ((rc = enqueue_describe(cam, cam->s.image, p, cam->s.kp_detect, cam->s.counters + 1))) return rc;
  return okc_detect_describe_finish(cam, keypoints, *n, n, descriptors);
}

int okc_detect_describe_enqueue(okc_camera* cam, const uint8_t* image, const okc_detect_params* dp,
                                const okc_describe_params* xp) {
  DetectArgs a;
  int rc = check_detect_params(cam, dp, &a);
  if (rc) return rc;
  if (!image || !xp) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  if ((rc = ensure_occupancy(cam, a))) return rc;
  if ((rc = upload_image(cam, image))) return rc;
  if ((rc = enqueue_detect(cam, cam->s.image, a))) return rc;
  if ((rc = enqueue_describe(cam, cam->s.image, xp, cam->s.kp_detect, cam->s.counters + 1))) return rc;
  return OKC_OK;
}

int okc_detect_describe_finish(okc_camera* cam, okc_keypoint* keypoints, int capacity, int* n, uint8_t* descriptors) {
  if (!cam || !keypoints || !n || !descriptors) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_counters, cam->s.counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_kp, cam->s.kp_desc, (size_t)cam->max_kp * sizeof(okc_keypoint),
                               cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_desc, cam->s.desc, (size_t)cam->max_kp * OKC_DESC_BYTES, cudaMemcpyDeviceToHost,
                               cam->stream));
  OKC_CUDA_TRY(cudaStreamSynchronize(cam->stream));
  int rc = check_flags(cam);
  if (rc) return rc;
  const int cnt = cam->h_counters[2];
  if (cnt > capacity) {
    set_error("%d keypoints exceed the caller's capacity %d", cnt, capacity);
    return OKC_ERR_CAPACITY;
  }
  memcpy(keypoints, cam->h_kp, (size_t)cnt * sizeof(okc_keypoint));
  memcpy(descriptors, cam->h_desc, (size_t)cnt * OKC_DESC_BYTES);
  *n = cnt;
  return OKC_OK;
}

int okc_detect_describe(okc_camera* cam, const uint8_t* image, const okc_detect_params* dp,
                        const okc_describe_params* xp, okc_keypoint* keypoints, int capacity, int* n,
                        uint8_t* descriptors) {
  int rc = okc_detect_describe_enqueue(cam, image, dp, xp);
  if (rc) return rc;
  return okc_detect_describe_finish(cam, keypoints, capacity, n, descriptors);
}

int okc_debug_scores(okc_camera* cam, const uint8_t* image, int32_t* scores) {
  if (!cam || !image || !scores) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  int rc;
  if ((rc = upload_image(cam, image))) return rc;
  DetectArgs a{};
  a.width = cam->W;
  a.height = cam->H;
  a.abs_threshold = 0x7fffffff;
  OKC_CUDA_TRY(cudaMemsetAsync(cam->s.counters, 0, 8 * sizeof(int), cam->stream));
  launch_harris_nms(cam->s.image, cam->s.scores, cam->s.bitmap, a, cam->stream);
  OKC_CUDA_TRY(cudaMemcpyAsync(scores, cam->s.scores, (size_t)cam->W * cam->H * sizeof(int32_t), cudaMemcpyDeviceToHost,
                               cam->stream));
  OKC_CUDA_TRY(cudaStreamSynchronize(cam->stream));
  return OKC_OK;
}

int okc_debug_candidates(okc_camera* cam, const uint8_t* image, int absolute_threshold, void* candidates, int cap,
                         int* n) {
  if (!cam || !image || !candidates || !n) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  int rc;
  if ((rc = upload_image(cam, image))) return rc;
  DetectArgs a{};
  a.width = cam->W;
  a.height = cam->H;
  a.abs_threshold = absolute_threshold;
  OKC_CUDA_TRY(cudaMemsetAsync(cam->s.counters, 0, 8 * sizeof(int), cam->stream));
  launch_harris_nms(cam->s.image, cam->s.scores, cam->s.bitmap, a, cam->stream);
  launch_order_candidates(cam->s.scores, cam->s.bitmap, cam->s.cand, cam->s.counters, cam->s.order_scratch, a,
                          cam->stream);
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->h_counters, cam->s.counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaStreamSynchronize(cam->stream));
  int cnt = cam->h_counters[0];
  if (cam->h_counters[3] == 1 || cnt > cap) {
    set_error("okc_debug_candidates: %d candidates exceed capacity", cnt);
    return OKC_ERR_CAPACITY;
  }
  // device order = the reference's std::sort order (no host sort here)
  uint64_t* keys = new uint64_t[cnt > 0 ? cnt : 1];
  OKC_CUDA_TRY(cudaMemcpy(keys, cam->s.cand, (size_t)cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  struct Pws { int32_t score; uint16_t x, y; };
  Pws* out = static_cast<Pws*>(candidates);
  for (int i = 0; i < cnt; ++i) {
    const uint32_t raster = (uint32_t)keys[i];
    out[i].score = cand_score(keys[i]);
    out[i].x = (uint16_t)(raster % cam->W);
    out[i].y = (uint16_t)(raster / cam->W);
  }
  delete[] keys;
  *n = cnt;
  return OKC_OK;
}

// Test hook for the std::sort reproduction alone: sorts n (<= W*H) arbitrary scores given in "raster order" and
// returns the permutation (order[i] = input index of the i-th output element).
int okc_debug_std_sort(okc_camera* cam, const int32_t* in_scores, int n, uint32_t* order) {
  if (!cam || !in_scores || !order || n < 0 || (size_t)n > (size_t)cam->W * cam->H || n > kCandCap) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(cam->ctx->device));
  const int wpr = (cam->W + 31) / 32;
  std::vector<uint32_t> bm((size_t)cam->H * wpr, 0u);
  for (int r = 0; r < n; ++r) {
    const int y = r / cam->W, x = r % cam->W;
    bm[(size_t)y * wpr + (x >> 5)] |= 1u << (x & 31);
  }
  DetectArgs a{};
  a.width = cam->W;
  a.height = cam->H;
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->s.bitmap, bm.data(), bm.size() * 4, cudaMemcpyHostToDevice, cam->stream));
  OKC_CUDA_TRY(cudaMemcpyAsync(cam->s.scores, in_scores, (size_t)n * 4, cudaMemcpyHostToDevice, cam->stream));
  OKC_CUDA_TRY(cudaMemsetAsync(cam->s.counters, 0, 8 * sizeof(int), cam->stream));
  launch_order_candidates(cam->s.scores, cam->s.bitmap, cam->s.cand, cam->s.counters, cam->s.order_scratch, a,
                          cam->stream);
  OKC_CUDA_TRY(cudaGetLastError());
  std::vector<uint64_t> keys((size_t)n + 1);
  OKC_CUDA_TRY(cudaMemcpyAsync(keys.data(), cam->s.cand, (size_t)n * 8, cudaMemcpyDeviceToHost, cam->stream));
  OKC_CUDA_TRY(cudaStreamSynchronize(cam->stream));
  for (int i = 0; i < n; ++i) order[i] = (uint32_t)keys[i];
  return OKC_OK;
}

}  // extern "C"

#include "capi_match.inc"
#include "capi_frame.inc"
