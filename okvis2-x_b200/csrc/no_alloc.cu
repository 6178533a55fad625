// okc_match_device_no_alloc: the plain best-match sweep (best[q] = the map row with the smallest Hamming distance
// strictly below the threshold, first row wins ties - what okc_match_batched / okc_match_sets compute) for callers
// that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// Two kernels, no atomics:
//   1. partial: grid (chunks of map rows, tiles of 128 queries); a thread keeps one query in registers, the CTA walks
//      its chunk's rows through a shared-memory tile; the thread's running best over the chunk goes to
//      scratch[chunk][q] as the key (dist << 32) | (row + index_offset), OKC_NO_MATCH when no row is below the threshold.
//   2. reduce: best[q] = min over the chunks (keys order by distance, then by global row = the first-wins rule).
// The scratch is a file-scope device array of a fixed size, so nothing is allocated per call or ever: every GPU that
// loads the library holds kScratchBytes (2 MiB) of it until the process exits.
//
// Every call shares that one array: calls that overlap on the GPU (different streams, or the same call captured into
// graphs that run concurrently) race on it. Nothing in it has to start at zero: kernel 1 overwrites every cell that
// kernel 2 of the same call reads (all chunks x nq), so the contents a previous call left behind are never read.
#include <cstdint>
#include <mutex>

#include <cuda_runtime.h>

#include "../../include/okvis_cuda.h"
#include "common.cuh"

namespace {

constexpr int kMaxQueries = OKC_NO_ALLOC_MAX_QUERIES;   // 4096
constexpr int kMaxChunks = 64;
constexpr size_t kScratchBytes = (size_t)kMaxQueries * kMaxChunks * sizeof(unsigned long long);
static_assert(kScratchBytes == (size_t)2 << 20, "the stated size of the fixed scratch: 2 MiB per GPU");
constexpr int kThreads = 128;      // queries per CTA
constexpr int kTileRows = 64;      // map rows staged per step (3 KB)
constexpr int kMinChunkRows = 512; // a chunk shorter than this is not worth a CTA

// In HOST code this name is only a handle for cudaGetSymbolAddress (below), never a device address.
static __device__ __align__(256) unsigned long long g_scratch[(size_t)kMaxQueries * kMaxChunks];

__global__ void __launch_bounds__(kThreads)
partial_best_kernel(const uint4* __restrict__ queries, int nq, const uint4* __restrict__ map, long long nm,
                    long long rowsPerChunk, unsigned threshold, unsigned long long indexOffset,
                    unsigned long long* __restrict__ scratch) {
  __shared__ uint4 s_rows[kTileRows * 3];
  const int q = blockIdx.y * kThreads + threadIdx.x;
  const int qSafe = q < nq ? q : nq - 1;
  unsigned w[12];
  {
    const uint4 a = __ldg(queries + (size_t)qSafe * 3), b = __ldg(queries + (size_t)qSafe * 3 + 1),
                c = __ldg(queries + (size_t)qSafe * 3 + 2);
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    w[8] = c.x; w[9] = c.y; w[10] = c.z; w[11] = c.w;
  }
  const long long lo = (long long)blockIdx.x * rowsPerChunk;
  const long long hi = lo + rowsPerChunk < nm ? lo + rowsPerChunk : nm;
  unsigned bestDist = threshold;          // strict '<' against the running best, which starts at the threshold
  unsigned long long bestKey = OKC_NO_MATCH;
  for (long long base = lo; base < hi; base += kTileRows) {
    const int rows = (int)(hi - base < kTileRows ? hi - base : kTileRows);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * 3; i += kThreads) s_rows[i] = __ldg(map + (size_t)base * 3 + i);
    __syncthreads();
    for (int r = 0; r < rows; ++r) {
      const uint4 a = s_rows[3 * r], b = s_rows[3 * r + 1], c = s_rows[3 * r + 2];
      const unsigned d = __popc(w[0] ^ a.x) + __popc(w[1] ^ a.y) + __popc(w[2] ^ a.z) + __popc(w[3] ^ a.w) +
                         __popc(w[4] ^ b.x) + __popc(w[5] ^ b.y) + __popc(w[6] ^ b.z) + __popc(w[7] ^ b.w) +
                         __popc(w[8] ^ c.x) + __popc(w[9] ^ c.y) + __popc(w[10] ^ c.z) + __popc(w[11] ^ c.w);
      if (d < bestDist) {
        bestDist = d;
        bestKey = ((unsigned long long)d << 32) | (unsigned long long)((unsigned long long)(base + r) + indexOffset);
      }
    }
  }
  if (q < nq) scratch[(size_t)blockIdx.x * kMaxQueries + q] = bestKey;
}

__global__ void reduce_best_kernel(const unsigned long long* __restrict__ scratch, int nq, int chunks,
                                   unsigned long long* __restrict__ best) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  unsigned long long k = OKC_NO_MATCH;
  for (int c = 0; c < chunks; ++c) {
    const unsigned long long v = scratch[(size_t)c * kMaxQueries + q];
    k = v < k ? v : k;
  }
  best[q] = k;
}

// Set-up, once per device, with that device current: the device address of the scratch array.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned long long* g_scratch_ptr[kMaxDevices] = {};

int scratch_of_current_device(unsigned long long** out) {
  int dev = 0;
  OKC_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    okc::set_error("okc_match_device_no_alloc: device ordinal %d is not supported (< %d)", dev, kMaxDevices);
    return OKC_ERR_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    OKC_CUDA_TRY(cudaGetSymbolAddress(&p, g_scratch));
    g_scratch_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_scratch_ptr[dev];
  return OKC_OK;
}

}  // namespace

extern "C" {

int okc_match_device_no_alloc(int device, const void* queries_dev, int nq, const void* map_dev, int64_t nm,
                              uint32_t threshold, uint64_t index_offset, uint64_t* best_dev, void* stream_handle) {
  if (nq < 0 || nm < 0 || (nq > 0 && (!queries_dev || !best_dev)) || (nm > 0 && !map_dev)) {
    okc::set_error("okc_match_device_no_alloc: null pointer or negative size");
    return OKC_ERR_INVALID;
  }
  if (nq > kMaxQueries) {
    okc::set_error("okc_match_device_no_alloc: %d queries exceed the fixed scratch (OKC_NO_ALLOC_MAX_QUERIES = %d); "
                   "split the queries or use okc_match_sets", nq, kMaxQueries);
    return OKC_ERR_CAPACITY;
  }
  if (nm > OKC_NO_ALLOC_MAX_MAP_ROWS) {
    okc::set_error("okc_match_device_no_alloc: %lld map rows exceed OKC_NO_ALLOC_MAX_MAP_ROWS (%lld)", (long long)nm,
                   (long long)OKC_NO_ALLOC_MAX_MAP_ROWS);
    return OKC_ERR_CAPACITY;
  }
  if ((reinterpret_cast<uintptr_t>(queries_dev) | reinterpret_cast<uintptr_t>(map_dev)) & 15) {
    okc::set_error("okc_match_device_no_alloc: descriptor rows must be 16-byte aligned");
    return OKC_ERR_INVALID;
  }
  if (device >= 0) OKC_CUDA_TRY(cudaSetDevice(device));
  unsigned long long* scratch = nullptr;
  int rc = scratch_of_current_device(&scratch);   // (first call per device: do it once outside a graph capture)
  if (rc) return rc;
  if (nq == 0) return OKC_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream_handle);
  long long chunks = (nm + kMinChunkRows - 1) / kMinChunkRows;
  if (chunks > kMaxChunks) chunks = kMaxChunks;
  if (chunks < 1) chunks = 1;
  long long rowsPerChunk = (nm + chunks - 1) / chunks;
  if (rowsPerChunk < 1) rowsPerChunk = 1;
  chunks = nm > 0 ? (nm + rowsPerChunk - 1) / rowsPerChunk : 1;
  const dim3 grid((unsigned)chunks, (unsigned)((nq + kThreads - 1) / kThreads));
  partial_best_kernel<<<grid, kThreads, 0, s>>>(static_cast<const uint4*>(queries_dev), nq,
                                                static_cast<const uint4*>(map_dev), (long long)nm, rowsPerChunk, threshold,
                                                (unsigned long long)index_offset, scratch);
  reduce_best_kernel<<<(nq + 255) / 256, 256, 0, s>>>(scratch, nq, (int)chunks,
                                                      reinterpret_cast<unsigned long long*>(best_dev));
  OKC_CUDA_TRY(cudaGetLastError());
  return OKC_OK;
}

}  // extern "C"
