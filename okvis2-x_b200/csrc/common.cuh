// okvis_cuda internal declarations shared by the .cu files.
// Everything is compiled with --fmad=false: the reference (brisk, -O3 -mssse3, no FMA target, no
// fast-math) rounds every float/double operation separately, and bit-exact parity needs the same here.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/okvis_cuda.h"

namespace okc {

constexpr int kDescWords = 12;         // 48 bytes = 12 x u32 = 3 x uint4
constexpr int kCandCap = 131072;       // candidate capacity per image (power of two, for the bitonic network)
constexpr int kNRot = 1024;            // brisk n_rot_
constexpr int kPoints = 66;
constexpr int kShortPairs = 384;
constexpr int kLongPairs = 856;

// Sort key of a candidate: high word orders by DEscending signed score, low word is the raster index
// (y * width + x), so an ascending u64 sort yields "score desc, raster asc".
__host__ __device__ inline uint64_t cand_key(int score, uint32_t raster) {
  return (static_cast<uint64_t>(~(static_cast<uint32_t>(score) ^ 0x80000000u)) << 32) | raster;
}
__host__ __device__ inline int cand_score(uint64_t key) {
  return static_cast<int>((~static_cast<uint32_t>(key >> 32)) ^ 0x80000000u);
}

struct PatternPoint { float x, y, sigma; };
struct LongPair { uint8_t i, j; int16_t pad; int wdx, wdy; };

// Host-built tables for one scale index (pattern.cc), uploaded once per context.
struct PatternTables {
  PatternPoint* points;    // [kNRot][kPoints]
  uchar2* short_pairs;     // [384] (i, j)
  LongPair* long_pairs;    // [856]
  float* lut;              // [31*31] uniformity LUT, layer 0
  int border;              // sizeList_[basicscale]
  int basicscale;
};

struct DeviceImageScratch {
  uint8_t* image;          // H*W
  int32_t* scores;         // H*W
  uint64_t* cand;          // kCandCap keys (std::sort order) + max_keypoints kept keys
  uint32_t* bitmap;        // H * ceil(W/32) words: NMS survivors
  uint8_t* order_scratch;  // order_candidates_kernel arrays when the candidates do not fit in shared memory
  int* counters;           // [0]=candidate count, [1]=kept count, [2]=described count, [3]=overflow flag
  uint8_t* occupancy;      // global fallback for the uniformity occupancy image
  size_t occupancy_bytes;
  okc_keypoint* kp_detect; // max_keypoints
  okc_keypoint* kp_desc;   // max_keypoints (after border filter)
  uint8_t* desc;           // max_keypoints * 48
};

struct DetectArgs {
  int width, height;
  int abs_threshold;       // (int) of the double parameter, as Get2dMaxima receives it
  float scaling;           // (float)(15.0 / (float)radius)
  unsigned long long max_kpt;
  int max_out;             // capacity of kp_detect
};

struct DescribeArgs {
  int width, height;
  int rotation_invariant;
  int camera_aware;
  float dir[3];
  float inv_f;             // 1.0f / virtualFocalLength_
};

void launch_harris_nms(const uint8_t* image, int32_t* scores, uint32_t* bitmap, const DetectArgs& a, cudaStream_t s);
size_t order_scratch_bytes();
void launch_order_candidates(const int32_t* scores, const uint32_t* bitmap, uint64_t* cand, int* counters,
                             uint8_t* gscratch, const DetectArgs& a, cudaStream_t s);
void launch_select_keypoints(const int32_t* scores, uint64_t* cand, int* counters, uint8_t* occ_global,
                             const float* lut, okc_keypoint* out, const DetectArgs& a, cudaStream_t s);
void launch_describe(const uint8_t* image, const okc_keypoint* kp_in, const int* n_in, okc_keypoint* kp_out,
                     uint8_t* desc, int* counters, const float* rays, const float* jac, const PatternTables& t,
                     const DescribeArgs& a, int max_kp, cudaStream_t s);

void set_error(const char* fmt, ...);

}  // namespace okc

#define OKC_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      okc::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return OKC_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)
