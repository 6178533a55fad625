// SURVEY 8(f) row f4, the data formats either side of the 48-byte descriptor:
//  * the text form of a descriptor in a saved map, okvis_ceres/src/Component.cpp:427-438 (write: 48 x
//    `setfill('0') << setw(2) << hex << uint32_t(byte)` = 96 lower-case hex characters) and :228-236 (read: every
//    2-character field through `std::stringstream(field) >> std::hex >> uint32_t`, the low byte kept);
//  * DBoW2's cluster centre of a set of descriptors, FBrisk::meanValue (okvis_frontend/src/FBrisk.cpp:25-60): bit b of
//    the mean is set when more than size()/2 (integer division) of the descriptors have it set.
// Byte work: one thread per output word, everything coalesced; the host-buffer entry points copy in and out.
#include "common.cuh"

namespace okc {

// ---- descriptor -> 96 hex characters: a thread turns 4 descriptor bytes (one 32-bit word) into 8 characters ----
__global__ void descriptors_to_hex_kernel(const unsigned* __restrict__ desc, uint2* __restrict__ hex, long long words) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= words) return;
  const unsigned w = desc[i];
  auto digit = [](unsigned v) { return v < 10 ? '0' + v : 'a' + (v - 10); };
  unsigned lo = 0, hi = 0;
  // byte k of the word -> characters 2k (high nibble) and 2k + 1 (low nibble)
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const unsigned b0 = (w >> (8 * k)) & 0xff, b1 = (w >> (8 * (k + 2))) & 0xff;
    lo |= (digit(b0 >> 4) | (digit(b0 & 15) << 8)) << (16 * k);
    hi |= (digit(b1 >> 4) | (digit(b1 & 15) << 8)) << (16 * k);
  }
  hex[i] = make_uint2(lo, hi);
}

// What `std::stringstream(two chars) >> std::hex >> uint32_t` leaves in the low byte (libstdc++ num_get, base 16):
// an optional sign, an optional "0x"/"0X" (which can never be followed by a digit inside two characters), then hex
// digits up to the first other character; no digit at all -> failbit and the value 0 (C++11); a leading '-' negates
// modulo 2^32. Leading white space (skipws, "C" locale) is skipped; it cannot occur in a file the reference reads (the
// field is cut from a token read with operator>>) but the function is total over all 65 536 fields this way.
__device__ __forceinline__ int hex_value(unsigned c) {
  if (c - '0' < 10u) return (int)(c - '0');
  if (c - 'a' < 6u) return (int)(c - 'a' + 10);
  if (c - 'A' < 6u) return (int)(c - 'A' + 10);
  return -1;
}
__device__ __forceinline__ unsigned parse_hex_field(unsigned c0, unsigned c1) {
  if (c0 == ' ' || c0 - '\t' < 5u) {   // space, \t \n \v \f \r
    const int v = hex_value(c1);
    return v < 0 ? 0u : (unsigned)v;
  }
  if (c0 == '+' || c0 == '-') {
    const int v = hex_value(c1);
    if (v < 0) return 0;
    return c0 == '-' ? (unsigned)(-v) & 0xffu : (unsigned)v;
  }
  const int a = hex_value(c0);
  if (a < 0) return 0;
  const int b = hex_value(c1);
  return b < 0 ? (unsigned)a : (unsigned)(a * 16 + b);
}

__global__ void descriptors_from_hex_kernel(const uint2* __restrict__ hex, unsigned* __restrict__ desc, long long words) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= words) return;
  const uint2 h = hex[i];
  unsigned w = 0;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    w |= parse_hex_field((h.x >> (16 * k)) & 0xff, (h.x >> (16 * k + 8)) & 0xff) << (8 * k);
    w |= parse_hex_field((h.y >> (16 * k)) & 0xff, (h.y >> (16 * k + 8)) & 0xff) << (8 * (k + 2));
  }
  desc[i] = w;
}

// ---- FBrisk::meanValue of every group: a CTA of 384 threads per group; thread t counts bit t%32 of word t/32.
// The 12 words of a row are read by the 12 warps (one broadcast load per warp and row). ----
__global__ void __launch_bounds__(384) descriptor_mean_kernel(const unsigned* __restrict__ desc,
                                                                const int* __restrict__ group_start,
                                                                unsigned* __restrict__ mean) {
  const int g = blockIdx.x, word = threadIdx.x >> 5, bit = threadIdx.x & 31;
  const int begin = group_start[g], end = group_start[g + 1];
  unsigned long long count = 0;
  for (int r = begin; r < end; ++r) count += (desc[(size_t)r * 12 + word] >> bit) & 1u;
  const unsigned long long half = (unsigned long long)(end > begin ? end - begin : 0) / 2;
  const unsigned m = __ballot_sync(0xffffffffu, count > half);
  if (bit == 0) mean[(size_t)g * 12 + word] = m;
}

}  // namespace okc

extern "C" {

namespace {
struct HostCall {   // one device block per call, the way geometry.cu's host-buffer entry points work
  uint8_t* d = nullptr;
  int rc = OKC_OK;
  ~HostCall() { cudaFree(d); }
  void check(cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  }
};
}  // namespace

int okc_descriptors_to_hex(okc_context* ctx, const uint8_t* descriptors, int n, char* hex) {
  if (!ctx || n < 0 || (n > 0 && (!descriptors || !hex))) return OKC_ERR_INVALID;
  if (n == 0) return OKC_OK;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  HostCall c;
  const size_t nb = (size_t)n * 48;
  OKC_CUDA_TRY(cudaMalloc(&c.d, nb * 3));
  c.check(cudaMemcpy(c.d, descriptors, nb, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (c.rc == OKC_OK) {
    const long long words = (long long)n * 12;
    okc::descriptors_to_hex_kernel<<<(unsigned)((words + 255) / 256), 256>>>(
        reinterpret_cast<const unsigned*>(c.d), reinterpret_cast<uint2*>(c.d + nb), words);
    c.check(cudaGetLastError(), "descriptors_to_hex_kernel");
    c.check(cudaMemcpy(hex, c.d + nb, nb * 2, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  return c.rc;
}

int okc_descriptors_from_hex(okc_context* ctx, const char* hex, int n, uint8_t* descriptors) {
  if (!ctx || n < 0 || (n > 0 && (!descriptors || !hex))) return OKC_ERR_INVALID;
  if (n == 0) return OKC_OK;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  HostCall c;
  const size_t nb = (size_t)n * 48;
  OKC_CUDA_TRY(cudaMalloc(&c.d, nb * 3));
  c.check(cudaMemcpy(c.d, hex, nb * 2, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (c.rc == OKC_OK) {
    const long long words = (long long)n * 12;
    okc::descriptors_from_hex_kernel<<<(unsigned)((words + 255) / 256), 256>>>(
        reinterpret_cast<const uint2*>(c.d), reinterpret_cast<unsigned*>(c.d + nb * 2), words);
    c.check(cudaGetLastError(), "descriptors_from_hex_kernel");
    c.check(cudaMemcpy(descriptors, c.d + nb * 2, nb, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  return c.rc;
}

int okc_descriptor_mean(okc_context* ctx, const uint8_t* descriptors, const int32_t* group_start, int n_groups,
                        uint8_t* mean) {
  if (!ctx || n_groups < 0 || (n_groups > 0 && (!group_start || !mean))) return OKC_ERR_INVALID;
  if (n_groups == 0) return OKC_OK;
  for (int g = 0; g < n_groups; ++g)
    if (group_start[g] < 0 || group_start[g + 1] < group_start[g]) {
      okc::set_error("okc_descriptor_mean: group_start must be non-negative and non-decreasing");
      return OKC_ERR_INVALID;
    }
  const size_t rows = (size_t)group_start[n_groups];
  if (rows > 0 && !descriptors) return OKC_ERR_INVALID;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  HostCall c;
  const size_t oG = (rows * 48 + 15) & ~(size_t)15, oM = oG + (((size_t)n_groups + 1) * 4 + 15 & ~(size_t)15);
  OKC_CUDA_TRY(cudaMalloc(&c.d, oM + (size_t)n_groups * 48));
  if (rows) c.check(cudaMemcpy(c.d, descriptors, rows * 48, cudaMemcpyHostToDevice), "cudaMemcpy");
  c.check(cudaMemcpy(c.d + oG, group_start, ((size_t)n_groups + 1) * 4, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (c.rc == OKC_OK) {
    okc::descriptor_mean_kernel<<<n_groups, 384>>>(reinterpret_cast<const unsigned*>(c.d),
                                                   reinterpret_cast<const int*>(c.d + oG),
                                                   reinterpret_cast<unsigned*>(c.d + oM));
    c.check(cudaGetLastError(), "descriptor_mean_kernel");
    c.check(cudaMemcpy(mean, c.d + oM, (size_t)n_groups * 48, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  return c.rc;
}

}  // extern "C"
