// okvis_cuda internal declarations shared by the .cu files.
// Everything is compiled with --fmad=false: the reference (brisk, -O3 -mssse3, no FMA target, no
// fast-math) rounds every float/double operation separately, and bit-exact parity needs the same here.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>

#include "../../include/okvis_cuda.h"
#include "../../include/okvis_cuda_debug.h"

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
  float* lut_upper;        // [31*31] uniformity LUT, layers >= 1 (octaves > 0)
  int border;              // sizeList_[basicscale]
  int basicscale;
};

constexpr int kMaxSlices = 64;
// ctrs[t] and the not-yet-divided reprErrors[t] of Frontend::matchToMapByThread (Frontend.cpp:1767, 1820-1821)
struct MatchStats { unsigned long long count; double sum; };

// One image of a batch ("slot"). Every kernel of the path is launched once per BATCH and indexes its slot with
// blockIdx (.z for the tiled kernels, .x for the one-CTA-per-image kernels), so that a multi-camera frame or several
// frames of a sequence fill the 148 SMs with one launch chain. The array lives in device memory; all pointers are
// fixed at batch creation except the ones marked (per enqueue), which the host refreshes with one small H2D copy.
struct Slot {
  // detect
  const uint8_t* image;    // (per enqueue) H*W u8: the slot's own upload buffer or a caller's device image
  int32_t* scores;         // H*W Harris scores
  uint32_t* bitmap;        // harris_bitmap_words(W, H) words: NMS survivors (layout: see harris_nms_kernel)
  uint64_t* cand;          // kCandCap keys (std::sort order) + max_keypoints kept keys
  uint8_t* order_scratch;  // order_candidates arrays when the candidates do not fit in shared memory
  uint8_t* occupancy;      // global fallback for the uniformity occupancy image
  int* counters;           // [0]=candidates [1]=kept [2]=described [3]=overflow flag [4]=match events
  okc_keypoint* kp_detect; // max_keypoints
  // describe
  okc_keypoint* kp_desc;   // max_keypoints (after the border filter)   } in the batch's output arena
  uint8_t* desc;           // max_keypoints * 48                        }
  const float* rays;       // camera-awareness maps of the slot's camera (or null)
  const float* jac;
  float inv_f;             // 1.0f / virtualFocalLength_
  int camera_aware;
  // map match (frame pipeline)
  const uint8_t* map;      // (per enqueue) descriptor pool rows
  const int* owner;        // (per enqueue) row -> landmark slot
  const double* proj;      // (per enqueue) landmark projections of this frame
  double thresh_sq;        // (per enqueue)
  int nm;                  // (per enqueue) rows in the pool (map == null: no match for this slot)
  int num_slices;          // (per enqueue) num_matching_threads: ctrs / reprErrors are defined per thread slice
  int event_cap;
  double* best_dist;       // max_keypoints   } output arena
  int* best_owner;         // max_keypoints   }
  MatchStats* stats;       // kMaxSlices      }
  unsigned long long* events;   // event_cap improvement events (k, row); consumed on the device by match_stats
  unsigned long long* sort_scratch;  // pow2(event_cap) keys, used when the events do not fit in shared memory
  float dir[3];            // (per enqueue) extraction direction (gravity in this camera's frame)
  // spatial bins of the landmark projections (bin_projections_kernel), rebuilt per enqueue: rows of the pool grouped
  // by the cell their projection falls into, ascending row index inside a cell
  int* bin_start;          // bins_x * bins_y + 1 offsets, then [ncells + 1] = 1 if a projection is not finite
  int* bin_rows;           // nm row indices
  double* bin_px;          // nm projections in bin order
  double* bin_py;
  int* bin_owner;          // nm landmark slots in bin order
  int bins_x, bins_y;      // 0: no bins for this slot (full scan)
  double bin_pad, bin_cell;   // cell(p) = floor((p + bin_pad) / bin_cell), bin_cell >= gate radius + 2 px = bin_pad
  // ---- frame pipeline beyond matchToMapByThread ----
  // back-projections of the described keypoints (Frame::computeBackProjections), written after describe when the
  // slot's camera has a geometry
  okc_pinhole pinhole;
  int has_geometry;
  double* e_C;             // max_keypoints x 3
  uint8_t* bp_valid;       // max_keypoints; counters[6] = number of valid ones
  // landmark pool packed on the device for this slot's pose (okc_batch_prepare_landmarks_enqueue): row count and the
  // per-row / per-landmark side data of the two map matchers
  const int* nm_dev;       // (per enqueue) rows in the pool, when it was packed on the device (else nm counts)
  const uint8_t* is3d;     // (per enqueue) per landmark: matchToMapByThread takes the rows of 3D landmarks, the
                           // uninitialised matcher the others; null: matchToMapByThread takes every row
  // pose of the slot's camera for this frame: T_WC = T_WS * T_SC
  double C_WC[9], r_WC[3]; // (per enqueue)
  double T_CW[12];         // (per enqueue) its inverse: rotation row-major, then translation
  double focal_mean;       // 0.5 * (focalLengthU + focalLengthV)
};

// One side of a 2D-2D match (Frontend::matchStereo / matchMotionStereo): a view's descriptors, keypoint sizes,
// back-projections and pose. Keypoint sizes come either as doubles (host-buffer entry points) or from the keypoints.
struct StereoSide {
  const uint8_t* desc;
  const double* size_d;     // or null:
  const okc_keypoint* kp;   //   (double)kp[k].size
  const double* e_C;
  const uint8_t* bp_valid;
  const int* n_dev;         // count still on the device (a batch slot), or null
  int n;                    // count, or the capacity when n_dev is set
  double C[9], r[3], f;
  const Slot* pose_slot;    // a batch slot: C / r are read from its record (refreshed every frame) instead
};
struct StereoJob {
  StereoSide v0, v1;
  const uint8_t* use0;      // motion stereo: eligibility of k0 (estimator state), or null = every k0
  // motion stereo against a batch slot: only the slot's keypoints without a landmark are candidates (the reference
  // packs them, Frontend.cpp:2034-2047): skipped unless both owners are < 0; null = every k1
  const int* v1_owner_a;
  const int* v1_owner_b;
  const int* v1_list;       // or: the candidates k1, ascending (a precompacted list of those keypoints), with
  const int* v1_list_n;     //   their number on the device
  double* dist;
  int* k1;
  double* hp;
  uint8_t* init;
  double* quality;          // motion stereo only
  // scratch of the two-kernel scheme (match.cu): per k0 up to kCandPerQuery candidates (k1 << 9) | dist and their count
  unsigned* cand;
  int* cand_count;
};

struct UninitArgs {
  double r_WC[3];
  double cos26Sigma, cos6Sigma;   // sigma = 1 / focalLength; cos(2.6 sigma), cos(6 sigma) from the host's libm
  double hammingThreshold;
};
// Frontend::matchToMapByThreadUnitialised for one camera image.
struct UninitJob {
  const uint8_t* queries;
  const double* kp_e_W;     // world rays of the keypoints, or null: C_WC * e_C.normalized() from
  const double* e_C;        //   the back-projections
  const uint8_t* bp_valid;  //   and their validity (Frontend.cpp:1862-1872)
  double C_WC[9];
  const Slot* pose_slot;    // a batch slot: C_WC and a.r_WC are read from its record (refreshed every frame) instead
  const uint8_t* use;       // eligibility flags, or null: a keypoint is eligible when
  const int* owner14;       //   matchToMapByThread left it without a landmark (owner14[k] < 0; null: always)
  const int* previousOwner; // loop-closure mode only, or null
  const int* nk_dev;
  int nk;
  const uint8_t* map;
  const int* owner;
  const uint8_t* is3d;      // rows of 3D landmarks are skipped (Frontend.cpp:1878-1880); null: no row is
  const double* map_e_W;
  const double* map_r_W;
  const int* nm_dev;
  int nm;
  const int* row_list;      // the rows to sweep, ascending (a precompacted list of the non-3D rows), or null: all rows
  const int* n_list_dev;    //   and their number
  const uint8_t* row_desc;  //   and, optionally, their descriptors in list order (contiguous: coalesced tile loads)
  int num_slices;
  UninitArgs a;
  double* dist;
  int* best;
  double* hp;
  uint8_t* hp_written;
  unsigned* ctr;                  // per keypoint, or null
  unsigned long long* slice_ctrs; // per thread slice (added to), or null
  // scratch of the two-kernel scheme (match.cu): per keypoint up to kCandPerQuery candidates (row << 9) | dist, count
  unsigned* cand;
  int* cand_count;
};
constexpr int kCandPerQuery = 32;   // candidate rows kept per query by the filter kernels of the gated matchers
void launch_match_stereo_jobs(const StereoJob* jobs, int nJobs, int max0, int max1, bool motion, double hammingThreshold,
                              int* cand_counts, size_t n_counts, cudaStream_t s);
void launch_match_uninit_jobs(const UninitJob* jobs, int nJobs, int maxNk, int maxRows, double hammingThreshold,
                              int* cand_counts, size_t n_counts, cudaStream_t s);
void launch_back_project_slots(const Slot* slots, int n, int max_kp, cudaStream_t s);
// per slot: the keypoints without a landmark after the two map matchers (owner_a[k] < 0 and owner_b[k] < 0; a null
// array counts as "none matched"), ascending, into list[slot * max_kp ..] with their number in count[slot]
struct UnmatchedArgs { const int* owner_a; const int* owner_b; const int* nk_dev; int* list; int* count; };
void launch_unmatched_lists(const UnmatchedArgs* args, int n, int max_kp, cudaStream_t s);
// exchange records of camera-sharded multiframes: count | keypoints | descriptors | back-projections | validity of the
// slots listed in slot_index[0 .. n), live rows only, into consecutive records of record_bytes
struct PackLayout { size_t off_n, off_kp, off_desc, off_eC, off_valid, bytes; };
void launch_pack_views(const Slot* slots, const int* slot_index, int n, int max_kp, PackLayout layout, uint8_t* dst,
                       cudaStream_t s);

// Landmark preparation of Frontend::matchToMap for one camera image (Frontend.cpp:1376-1500): inputs of the shared
// landmark pool, the pose of this image, and where its pool is packed.
struct PrepareArgs {
  const double* hp_W;          // L x 4
  const double* quality;       // L
  const uint8_t* visible;      // L (okc_project_landmarks)
  const long long* obs_offset; // L + 1
  const int* obs_view;         // per observation: view (frame, camera) index, in the reference's iteration order
  const int* obs_kp;           // per observation: keypoint index inside the view
  const double* view_C;        // V x 9, T_WC_old.C() row-major
  const double* view_r;        // V x 3, T_WC_old.r()
  const long long* view_row0;  // V: first row of the view in the descriptor / back-projection pools
  const uint8_t* desc;         // N x 48
  const double* e_C;           // N x 3 (Frame::backProjections_)
  double r_WC1[3];             // T_WC1.r() of the camera being matched
  double focal_sum;            // focalLengthU() + focalLengthV() (Frontend.cpp:1361-1363)
  int loop_closure;            // loopClosureLandmarksToUseExclusively != nullptr
  int L;
  // per landmark, 3 slots
  uint8_t* is3d;
  int* rows;                   // o + 1 for visible landmarks, 0 otherwise
  int* slot_obs;               // L x 3: observation index that filled the slot (-1: never written)
};
struct PrepSlot {
  PrepareArgs a;               // (a.visible = visible below)
  okc_pinhole cam;
  double T_CW[12];
  const Slot* pose_slot;       // a batch slot: T_CW and a.r_WC1 are read from its record (refreshed every frame) instead
  double repr_threshold;
  double* proj_xy;             // L x 2
  uint8_t* status;             // L
  uint8_t* visible;            // L
  int* row_offset;             // L + 1; [L] = rows in the pool
  uint8_t* pool;               // 3 L x 48
  int* owner;                  // 3 L
  double* e_W;                 // 3 L x 3
  double* r_W;                 // 3 L x 3
  uint8_t* filled;             // 3 L
  int* row_obs;                // 3 L
  int* u_count;                // L     scratch
  int* u_offset;               // L + 1 rows of the non-3D landmarks before landmark l; [L] = their number
  int* u_rows;                 // 3 L   those rows (indices into the pool), ascending
  uint8_t* u_desc;             // 3 L x 48: their descriptors, contiguous in list order (the filter kernel's rows)
};
void launch_prepare_landmarks_slots(const PrepSlot* prep, int n, int L, cudaStream_t s);
int check_pinhole(const okc_pinhole* cam);   // OKC_OK or OKC_ERR_INVALID (with the error string set)


struct DetectArgs {
  int width, height;
  int abs_threshold;       // (int) of the double parameter, as Get2dMaxima receives it
  float scaling;           // (float)(15.0 / (float)radius)
  unsigned long long max_kpt;
  int max_out;             // capacity of kp_detect
  // keypoint coordinates of the layer: x = kp_scale * ((x + dx) + kp_offset) (scale-space-layer-inl.h:412-419)
  double kp_scale = 1.0, kp_offset = 0.0;
  int octave = 0;          // _layerNumber / 2
  bool dense_scores = false;   // write the whole score map (scale-space layers, okc_debug_scores), not only the maxima
};

struct DescribeArgs {
  int width, height;
  int rotation_invariant;
};

// All launchers: `slots` = device array, `n` = number of slots in this launch.
// aligned: every slot's image pointer is 4-byte aligned and width % 4 == 0 (32-bit pixel loads, int4 score stores)
void launch_harris_nms(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s, bool aligned);
size_t harris_score_ints(int W, int H);
size_t harris_bitmap_words(int W, int H);
void harris_bitmap_pos(int W, int x, int y, size_t* word, int* bit);
size_t order_scratch_bytes();
void launch_order_candidates(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s);
void launch_select_keypoints(const Slot* slots, int n, const float* lut, const DetectArgs& a, cudaStream_t s);
size_t select_scratch_bytes(const DetectArgs& a);   // per-slot Slot::occupancy bytes
// kp_from_detect: describe the keypoints the detector left in kp_detect (count = counters[1])
void launch_describe(const Slot* slots, int n, const PatternTables& t, const DescribeArgs& a, int max_kp,
                     cudaStream_t s);
void launch_match_to_map_slots(const Slot* slots, int n, int max_kp, double hammingThreshold, cudaStream_t s);
void launch_match_stats(const Slot* slots, int n, cudaStream_t s);
void launch_bin_projections(const Slot* slots, int n, cudaStream_t s);
void launch_match_to_map_binned(const Slot* slots, int n, int max_kp, double hammingThreshold, cudaStream_t s);
constexpr int kMaxBinCells = 4096;
// scale space (octaves > 0)
void launch_halfsample8(const uint8_t* src, int sw, int sh, uint8_t* dst, cudaStream_t s);
void launch_twothirdsample8(const uint8_t* src, int sw, int sh, uint8_t* dst, cudaStream_t s);
void launch_scale_nms(const int32_t* scores, uint32_t* bitmap, int W, int H, const int32_t* aboveScores, int aW,
                      int aH, double aScale, double aOffset, int aStep, const int32_t* belowScores, int bW, int bH,
                      double bScale, double bOffset, int bStep, cudaStream_t s);
void launch_concat_layers(const Slot* layers, int nLayers, const Slot& out, int maxOut, cudaStream_t s);

void set_error(const char* fmt, ...);

// Test hooks (include/okvis_cuda_debug.h): A/B baselines of the tests and forced launch shapes. Process-wide and read
// at launch time; a test sets them while no other call is in flight. Production callers never touch them.
struct Knobs {
  std::atomic<int> select_sequential{0};  // 1: one-candidate-at-a-time selection kernel instead of the wave kernel
  std::atomic<int> match_plain{0};        // 1: plain 12-word Hamming sweep instead of the lower-bound kernel
  std::atomic<int> match_tile_scan{0};    // 1: frame-pipeline map match scans every row instead of the spatial bins
  std::atomic<int> harris_rows{0};        // > 0: image rows one Harris warp walks (0: cost model)
};
extern Knobs g_knobs;

}  // namespace okc

// Every allocation and release the library makes goes through these wrappers (the macros below rewrite the call
// sites), so that a test can show that the per-frame entry points allocate nothing once a batch is warm:
// okc_debug_alloc_counts (include/okvis_cuda_debug.h) reads the counters.
namespace okc {
extern std::atomic<unsigned long long> g_alloc_calls, g_free_calls;
template <typename T>
inline cudaError_t counted_malloc(T** p, size_t n) {
  g_alloc_calls.fetch_add(1, std::memory_order_relaxed);
  return ::cudaMalloc(reinterpret_cast<void**>(p), n);
}
template <typename T>
inline cudaError_t counted_malloc_host(T** p, size_t n) {
  g_alloc_calls.fetch_add(1, std::memory_order_relaxed);
  return ::cudaMallocHost(reinterpret_cast<void**>(p), n);
}
inline cudaError_t counted_free(void* p) {
  g_free_calls.fetch_add(1, std::memory_order_relaxed);
  return ::cudaFree(p);
}
inline cudaError_t counted_free_host(void* p) {
  g_free_calls.fetch_add(1, std::memory_order_relaxed);
  return ::cudaFreeHost(p);
}
}  // namespace okc
#define cudaMalloc(p, n) okc::counted_malloc((p), (n))
#define cudaMallocHost(p, n) okc::counted_malloc_host((p), (n))
#define cudaFree(p) okc::counted_free((p))
#define cudaFreeHost(p) okc::counted_free_host((p))

#define OKC_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      okc::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return OKC_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)
