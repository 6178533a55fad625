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

struct okc_descriptor_set {
  okc_context* ctx = nullptr;
  void* rows = nullptr;
  int64_t n = 0;
};

namespace okc {
int launch_hamming_best(const void* queries, int nq, const void* map, long long nm, unsigned threshold,
                        unsigned long long indexOffset, unsigned long long* best, cudaStream_t s,
                        unsigned long long noMatch = ~0ull);
}  // namespace okc

namespace okc {
Knobs g_knobs;
std::atomic<unsigned long long> g_alloc_calls{0}, g_free_calls{0};
}

extern "C" {

const char* okc_last_error_string(void) { return g_err; }
int okc_version(void) { return 100; }

/* Test hooks, declared in include/okvis_cuda_debug.h. */
int okc_debug_set_select_wave(int variant) {   /* 0 = wave kernel (default), 1 = sequential-visit kernel */
  if (variant != 0 && variant != 1) return OKC_ERR_INVALID;
  okc::g_knobs.select_sequential = variant;
  return OKC_OK;
}
int okc_debug_set_harris_rows(int rows) {      /* 0 = chosen per launch by the cost model */
  if (rows < 0 || rows > 4096) return OKC_ERR_INVALID;
  okc::g_knobs.harris_rows = rows;
  return OKC_OK;
}
int okc_debug_set_match_prefix(int words) {    /* 0 = lower-bound sweep kernel (default), 12 = plain 12-word sweep */
  if (words != 0 && words != 12) return OKC_ERR_INVALID;
  okc::g_knobs.match_plain = words == 12 ? 1 : 0;
  return OKC_OK;
}
int okc_debug_set_match_binned(int on) {       /* 1 = spatial bins (default), 0 = scan of every map row */
  okc::g_knobs.match_tile_scan = on ? 0 : 1;
  return OKC_OK;
}

int okc_context_destroy(okc_context* c);

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
  auto build = [&]() -> int {
    OKC_CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    build_host_tables(&c->host);
    const HostTables& t = c->host;
    OKC_CUDA_TRY(cudaMalloc(&c->tables.points, t.points.size() * sizeof(HostPatternPoint)));
    OKC_CUDA_TRY(cudaMalloc(&c->tables.short_pairs, t.short_pairs.size()));
    OKC_CUDA_TRY(cudaMalloc(&c->tables.long_pairs, t.long_pairs.size() * sizeof(HostLongPair)));
    OKC_CUDA_TRY(cudaMalloc(&c->tables.lut, t.lut.size() * sizeof(float)));
    OKC_CUDA_TRY(cudaMalloc(&c->tables.lut_upper, t.lut_upper.size() * sizeof(float)));
    OKC_CUDA_TRY(cudaMemcpy(c->tables.lut_upper, t.lut_upper.data(), t.lut_upper.size() * sizeof(float),
                            cudaMemcpyHostToDevice));
    OKC_CUDA_TRY(cudaMemcpy(c->tables.points, t.points.data(), t.points.size() * sizeof(HostPatternPoint),
                            cudaMemcpyHostToDevice));
    OKC_CUDA_TRY(cudaMemcpy(c->tables.short_pairs, t.short_pairs.data(), t.short_pairs.size(), cudaMemcpyHostToDevice));
    OKC_CUDA_TRY(cudaMemcpy(c->tables.long_pairs, t.long_pairs.data(), t.long_pairs.size() * sizeof(HostLongPair),
                            cudaMemcpyHostToDevice));
    OKC_CUDA_TRY(cudaMemcpy(c->tables.lut, t.lut.data(), t.lut.size() * sizeof(float), cudaMemcpyHostToDevice));
    c->tables.border = t.border;
    c->tables.basicscale = t.basicscale;
    return OKC_OK;
  };
  const int rc = build();
  if (rc) {
    okc_context_destroy(c);   // frees whatever was allocated before the failure
    return rc;
  }
  *out = c;
  return OKC_OK;
}

int okc_context_destroy(okc_context* c) {
  if (!c) return OKC_OK;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  cudaFree(c->tables.points);
  cudaFree(c->tables.short_pairs);
  cudaFree(c->tables.long_pairs);
  cudaFree(c->tables.lut);
  cudaFree(c->tables.lut_upper);
  for (auto& b : c->d) b.release();
  for (auto& b : c->h) b.release();
  if (c->stream) cudaStreamDestroy(c->stream);
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


}  // extern "C"

#include "capi_match.inc"
#include "capi_batch.inc"
#include "capi_frame.inc"
#include "capi_pipeline.inc"
