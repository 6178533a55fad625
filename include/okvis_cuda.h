/* okvis_cuda - C ABI of the B200 (sm_100a) BRISK frontend hot path of OKVIS2-X.
 *
 * Plain C, opaque handles, int status return (0 = OK), no exceptions and no
 * torch / OpenCV / Eigen types across the boundary. Caller owns every host
 * buffer; the library copies in and out and keeps no caller pointer after a
 * call returns. There is NO CPU fallback: without a CUDA device every compute
 * entry point returns OKC_ERR_CUDA.
 *
 * Each entry point names the reference interface it replaces
 * (paths relative to the OKVIS2-X tree; "brisk/" = external/brisk/).
 */
#ifndef OKVIS_CUDA_H_
#define OKVIS_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library itself is built with -fvisibility=hidden */
#endif

enum {
  OKC_OK = 0,
  OKC_ERR_INVALID = 1,   /* bad argument; mirrors the reference's std::runtime_error cases */
  OKC_ERR_CUDA = 2,      /* CUDA runtime error or no device */
  OKC_ERR_CAPACITY = 3   /* an output buffer or internal scratch is too small */
};

/* Same 28-byte layout as cv::KeyPoint {Point2f pt; float size, angle, response; int octave, class_id}. */
typedef struct okc_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} okc_keypoint;

#define OKC_DESC_BYTES 48 /* 384-bit BRISK2 descriptor; brisk/src/brisk-descriptor-extractor.cc:291-297 */
#define OKC_NO_MATCH 0xffffffffffffffffull

typedef struct okc_context okc_context; /* one per process and device */
typedef struct okc_camera okc_camera;   /* one per camera: stream, scratch, maps (like the per-camera
                                           detector/extractor pair of Frontend.cpp:2637-2656) */

const char* okc_last_error_string(void);
int okc_version(void);

/* device < 0: use the current device. */
int okc_context_create(int device, okc_context** ctx);
int okc_context_destroy(okc_context* ctx);
int okc_context_device(const okc_context* ctx);

/* Replaces: brisk::BriskDescriptorExtractor::setCameraProperties (brisk/include/brisk/brisk-descriptor-extractor.h:112-118)
 * as called from Frontend::detectAndDescribe (okvis_frontend/src/Frontend.cpp:214-231).
 * rays: float[H*W*3] (CV_32FC3) or NULL; jacobians: float[H*W*6] (CV_32FC6) or NULL (both or neither).
 * max_keypoints bounds the per-image keypoint capacity (scratch is sized once). */
int okc_camera_create(okc_context* ctx, int width, int height, const float* rays, const float* jacobians,
                      float virtual_focal_length, int max_keypoints, okc_camera** cam);
int okc_camera_destroy(okc_camera* cam);

typedef struct okc_detect_params {
  double uniformity_radius;   /* FrontendParameters::detection_threshold -> ctor arg 1 (Frontend.cpp:2645-2648) */
  uint64_t octaves;           /* 0 in every shipped config; > 0 (scale space, 2*octaves layers, <= 8) is available
                                 through the single-camera entry points, not through okc_batch_* */
  double absolute_threshold;
  uint64_t max_num_keypoints;
} okc_detect_params;

typedef struct okc_describe_params {
  int rotation_invariant;     /* okvis: true  (Frontend.cpp:125) */
  int scale_invariant;        /* okvis: false (Frontend.cpp:126); true returns OKC_ERR_INVALID */
  float extraction_direction[3]; /* setExtractionDirection (Frontend.cpp:234-238); used when maps are set */
} okc_describe_params;

/* Replaces brisk::ScaleSpaceFeatureDetector<HarrisScoreCalculator>::detect
 * (brisk/include/brisk/scale-space-feature-detector.h:89-163) as called by okvis::Frame::detect
 * (okvis_cv/include/okvis/implementation/Frame.hpp:120-134).
 * image: u8, continuous rows (stride == width, as the reference demands, :95-97).
 * keypoints[capacity] receives *n keypoints in the reference order (descending score). */
int okc_detect(okc_camera* cam, const uint8_t* image, const okc_detect_params* p, okc_keypoint* keypoints,
               int capacity, int* n);

/* Replaces brisk::BriskDescriptorExtractor::compute / doDescriptorComputation
 * (brisk/src/brisk-descriptor-extractor.cc:575-805) as called by okvis::Frame::describe (Frame.hpp:140-155).
 * keypoints is in/out: border-filtered in order (n shrinks), angle written back. descriptors: u8[n*48]. */
int okc_describe(okc_camera* cam, const uint8_t* image, const okc_describe_params* p, okc_keypoint* keypoints,
                 int* n, uint8_t* descriptors);

/* Fused detect + describe of Frontend::detectAndDescribe -> MultiFrame::detect + describe
 * (Frontend.cpp:204-262): image uploaded once, keypoints never leave the device between the stages. */
int okc_detect_describe(okc_camera* cam, const uint8_t* image, const okc_detect_params* dp,
                        const okc_describe_params* xp, okc_keypoint* keypoints, int capacity, int* n,
                        uint8_t* descriptors);

/* Same, asynchronous halves for overlapping several cameras (one std::thread per camera in
 * ThreadedSlam.cpp:676-708 becomes one stream per camera): enqueue returns immediately,
 * finish waits for that camera's stream and copies the results out. */
int okc_detect_describe_enqueue(okc_camera* cam, const uint8_t* image, const okc_detect_params* dp,
                                const okc_describe_params* xp);
int okc_detect_describe_finish(okc_camera* cam, okc_keypoint* keypoints, int capacity, int* n,
                               uint8_t* descriptors);

/* Intermediate products, for parity tests (not used by the reference's callers):
 * scores int32[H*W] = brisk::HarrisScoresSSE (brisk/src/harris-scores.cc:65-301);
 * candidates {int32 score; uint16 x, y}[cap] = HarrisScoreCalculator::Get2dMaxima
 * (brisk/src/harris-score-calculator.cc:70-119) in the order std::sort leaves them at the top of
 * EnforceKeyPointUniformity (uniformity-enforcement-inl.h:67; libstdc++ introsort, tie order included). */
int okc_debug_scores(okc_camera* cam, const uint8_t* image, int32_t* scores);
int okc_debug_candidates(okc_camera* cam, const uint8_t* image, int absolute_threshold, void* candidates, int cap,
                         int* n);

/* ---- Matching. Replaces the loops around brisk::Hamming::PopcntofXORed(a, b, 3)
 * (brisk/include/brisk/internal/hamming-inl.h:95-144) in okvis_frontend/src/Frontend.cpp. ---- */

/* Plain brute force: for each query the map row with the smallest Hamming distance strictly below
 * `threshold`, first row wins ties (strict '<' running minimum). Frontend::verifyRecognisedPlace
 * (Frontend.cpp:316-345) and the 10k x 1M sweep. best[q] = (dist << 32) | (index + index_offset), or
 * OKC_NO_MATCH. Host buffers. */
int okc_match_batched(okc_context* ctx, const uint8_t* queries, int nq, const uint8_t* map, int64_t nm,
                      uint32_t threshold, uint64_t index_offset, uint64_t* best);

/* Device-resident variant for sharded maps: upload once, sweep many times. */
typedef struct okc_descriptor_set okc_descriptor_set;
int okc_descriptor_set_create(okc_context* ctx, const uint8_t* rows, int64_t n, okc_descriptor_set** set);
int okc_descriptor_set_destroy(okc_descriptor_set* set);
/* best_device != NULL: leave the result in this device buffer (uint64[nq]) for a following NCCL gather
 * and skip the copy to best_host (may be NULL). stream_handle: a cudaStream_t or 0. */
int okc_match_sets(okc_context* ctx, const okc_descriptor_set* queries, const okc_descriptor_set* map,
                   uint32_t threshold, uint64_t index_offset, uint64_t* best_host, void* best_device,
                   void* stream_handle);
/* The same into a device buffer of int64 keys with INT64_MAX (instead of OKC_NO_MATCH) for "no match": the keys of all
 * shards then merge with ONE NCCL all-reduce(min) - distance first, global row index second, i.e. the reference's
 * first-wins rule - without an epilogue kernel. */
int okc_match_sets_ordered(okc_context* ctx, const okc_descriptor_set* queries, const okc_descriptor_set* map,
                           uint32_t threshold, uint64_t index_offset, void* best_device_i64, void* stream_handle);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The same result as
 * okc_match_batched / okc_match_sets - best[q] = (dist << 32) | (row + index_offset) of the map row with the smallest
 * Hamming distance strictly below `threshold`, first row wins ties, OKC_NO_MATCH otherwise - with every buffer supplied
 * by the caller as DEVICE pointers (queries_dev: u8[nq][48], map_dev: u8[nm][48], both 16-byte aligned; best_dev:
 * uint64[nq]) and the kernels' scratch (per-chunk partial results) in a fixed device array inside the library: the call
 * allocates nothing - no cudaMalloc / cudaFree / allocating library routine - and only enqueues two kernels on
 * `stream_handle` (a cudaStream_t, or 0 for the default stream); it does not synchronise.
 * Limits: nq <= OKC_NO_ALLOC_MAX_QUERIES and nm <= OKC_NO_ALLOC_MAX_MAP_ROWS, beyond them OKC_ERR_CAPACITY. The scratch
 * (2 MiB) is held by every GPU that loads the library until the process exits, and is shared by ALL calls: calls that
 * overlap on the GPU (different streams, concurrent graph launches) race on it - serialise them. Nothing in it needs
 * clearing between calls (a call overwrites every cell it later reads). device < 0: the current device. The first call
 * on a device looks up the scratch address (cudaGetSymbolAddress): make one call (nq = 0 suffices) before a capture. */
#define OKC_NO_ALLOC_MAX_QUERIES 4096
#define OKC_NO_ALLOC_MAX_MAP_ROWS 2147483647LL
int okc_match_device_no_alloc(int device, const void* queries_dev, int nq, const void* map_dev, int64_t nm,
                              uint32_t threshold, uint64_t index_offset, uint64_t* best_dev, void* stream_handle);

/* 3D-2D map matching, Frontend::matchToMapByThread (Frontend.cpp:1744-1827) for all thread slices at once.
 * Keypoint k (descriptor queries[k], position kp_xy[2k..]) with use[k] != 0 is compared with every
 * descriptor row m (owner landmark slot map_owner[m], ascending = std::map order) whose landmark projection
 * proj_xy[2*owner] lies within sqrt(thresh_sq) pixels. Outputs per k: best_dist[k] (double, init = threshold),
 * best_owner[k] (landmark slot or -1). Per thread slice t (num_slices = num_matching_threads): ctrs[t] and
 * repr_errors[t] exactly as the reference accumulates them (improvement events in landmark-major order,
 * final division included; NaN when ctrs[t] == 0). */
int okc_match_to_map(okc_context* ctx, const uint8_t* queries, const double* kp_xy, const uint8_t* use, int nk,
                     const uint8_t* map, const int32_t* map_owner, int nm, const double* proj_xy, int num_landmarks,
                     double thresh_sq, double hamming_threshold, int num_slices, double* best_dist,
                     int32_t* best_owner, uint64_t* ctrs, double* repr_errors);

/* 2D-2D stereo matching, the k0/k1 loops of Frontend::matchStereo (Frontend.cpp:2259-2316) with
 * triangulation::triangulateFast (okvis_frontend/src/stereo_triangulation.cpp:30-111) on the device in FP64.
 * Per camera i: descriptors desc_i[n_i*48], keypoint sizes size_i[n_i], back-projections e_C_i[n_i*3] with
 * validity flags bp_valid_i[n_i], pose T_WC_i as row-major 3x3 rotation C_i[9] and position r_i[3], mean focal
 * length f_i. Outputs per k0: dist[k0] (double, = threshold when unmatched), k1_match[k0], hp_W[4*k0..]
 * (normalised homogeneous point), initialisable[k0]. */
typedef struct okc_stereo_view {
  const uint8_t* desc;
  const double* size;
  const double* e_C;
  const uint8_t* bp_valid;
  int n;
  double C_WC[9];
  double r_WC[3];
  double f;
} okc_stereo_view;
int okc_match_stereo(okc_context* ctx, const okc_stereo_view* v0, const okc_stereo_view* v1,
                     double hamming_threshold, double* dist, int32_t* k1_match, double* hp_W,
                     uint8_t* initialisable);

/* 2D-2D matching against an older keyframe, the worker lambda of Frontend::matchMotionStereo (Frontend.cpp:2053-2137).
 * v0 = older frame (all keypoints; use0[k0] != 0 where the estimator-state tests of :2056-2084 let k0 through),
 * v1 = the current frame's UNMATCHED keypoints packed contiguously as the reference packs desc1 (:2034-2047);
 * k1_match indexes that packed list (-1: none). Outputs per k0 as MatchInfo (:1959-1966) before the 4 px
 * reprojection check of :2140-2147, which needs the camera model and stays in the adapter: dist (= threshold when
 * unmatched), k1, hp_W (w = 1), initialisable, quality = acos(...). */
int okc_match_motion_stereo(okc_context* ctx, const okc_stereo_view* v0, const uint8_t* use0, const okc_stereo_view* v1,
                            double hamming_threshold, double* dist, int32_t* k1_match, double* hp_W,
                            uint8_t* initialisable, double* quality);

/* 3D-2D matching of not-yet-initialised landmarks, Frontend::matchToMapByThreadUnitialised
 * (Frontend.cpp:1830-1956): epipolar-plane, divergence, triangulateFast and 0.2 m gates on the device in FP64.
 * Per map row m: owner slot, ray e_W[3m..], origin r_W[3m..]. Per keypoint k: e_W (world ray), use flag and
 * previous landmark slot (or -1; loop-closure mode only, Frontend.cpp:1941-1944). */
int okc_match_to_map_uninitialised(okc_context* ctx, const uint8_t* queries, const double* kp_e_W,
                                   const uint8_t* use, const int32_t* previous_owner, int nk, const uint8_t* map,
                                   const int32_t* map_owner, const double* map_e_W, const double* map_r_W, int nm,
                                   const double r_WC[3], double focal_length, double hamming_threshold,
                                   int num_slices, double* best_dist, int32_t* best_owner, double* hps_W,
                                   uint8_t* hps_written, uint64_t* ctrs);

/* ---- Camera geometry next to the path (rows f2 and the first half of f1). FP64 on the device.
 * The Eigen evaluation order is restated, not pinned (Eigen is not available where this was built): results agree
 * with the CPU restatement in oracle/port bit for bit and with the reference up to that caveat. ---- */
#define OKC_DISTORTION_NONE 0        /* cameras::NoDistortion */
#define OKC_DISTORTION_RADTAN 1      /* cameras::RadialTangentialDistortion: k = {k1, k2, p1, p2} */
#define OKC_DISTORTION_EQUIDISTANT 2 /* cameras::EquidistantDistortion:      k = {k1, k2, k3, k4} */
typedef struct okc_pinhole {         /* cameras::PinholeCamera<D> intrinsics (okvis_cv/include/okvis/cameras/PinholeCamera.hpp) */
  int width, height;
  double fu, fv, cu, cv;
  int distortion;
  double k[8];
} okc_pinhole;
/* cameras::ProjectionStatus (okvis_cv/include/okvis/cameras/CameraBase.hpp:38-45) */
#define OKC_PROJ_SUCCESSFUL 0
#define OKC_PROJ_OUTSIDE_IMAGE 1
#define OKC_PROJ_MASKED 2
#define OKC_PROJ_BEHIND 3
#define OKC_PROJ_INVALID 4

/* okvis::Frame::computeBackProjections (okvis_cv/include/okvis/implementation/Frame.hpp:158-173): for every keypoint
 * PinholeCamera::backProject (PinholeCamera.hpp:566-585) = unscale, D::undistort (Gauss-Newton on D::distort),
 * direction (x, y, 1). e_C[3k..] receives the direction also when the iteration did not succeed (as the reference
 * stores it), valid[k] the success flag, *n_valid the return value of computeBackProjections. Host buffers. */
int okc_back_project(okc_context* ctx, const okc_pinhole* cam, const okc_keypoint* keypoints, int n, double* e_C,
                     uint8_t* valid, int* n_valid);

/* Field-of-view test of the landmark loop of Frontend::matchToMap (Frontend.cpp:1376-1407): hp_C = T_CW * hp_W
 * (T_CW = 3x3 rotation, row-major, followed by the translation: 12 doubles), projectHomogeneous, and the window
 * [-repr_threshold, width + repr_threshold] x [-repr_threshold, height + repr_threshold]. Per landmark:
 * proj_xy[2l..] (the projection that becomes LandmarkToMatch::projection and the proj_xy input of okc_match_to_map),
 * status[l] (OKC_PROJ_*), visible[l] = landmark survives the `continue`s of :1393-1405. Host buffers. */
int okc_project_landmarks(okc_context* ctx, const okc_pinhole* cam, const double T_CW[12], const double* hp_W, int n,
                          double repr_threshold, double* proj_xy, uint8_t* status, uint8_t* visible);

/* Keyframe-overlap masks of Frontend::doWeNeedANewKeyframe (okvis_frontend/src/Frontend.cpp:1186-1295), SURVEY 8(f)
 * row f3: per view (one camera image of one multiframe, width[v] x height[v]) the H/10 x W/10 "detections" mask (a
 * filled cv::circle of radius int(min(rows, cols) * 0.09) at cv::Point(pt * 0.1) for every keypoint) and "matches"
 * mask (the same for keypoints with matched[k] != 0: landmarkId != 0 for the current frame, landmarkId in the
 * current frame's id set for the other frames - decided by the caller), and cv::countNonZero of their AND
 * (intersection[v]) and OR (uni[v]). Keypoints of view v = keypoints[kp_offset[v] .. kp_offset[v+1]). The caller
 * sums the counts over the cameras of a multiframe and forms the ratios (:1233, :1280). Host buffers. */
int okc_keyframe_overlap(okc_context* ctx, int n_views, const int* width, const int* height, const int64_t* kp_offset,
                         const okc_keypoint* keypoints, const uint8_t* matched, int32_t* intersection, int32_t* uni);

/* Landmark preparation of Frontend::matchToMap, second half (Frontend.cpp:1409-1500), SURVEY 8(f) row f1: for every
 * landmark with visible[l] != 0 (okc_project_landmarks) walk its observations obs_offset[l] .. obs_offset[l+1] in the
 * order given (the reference iterates the observation set in reverse: newest first), set is3d[l] (:1433-1439), drop
 * observations beyond 0.6 rad viewpoint change or 50 % scale change (not in loop-closure mode), keep the best three by
 * score with the reference's replace-the-worst-slot rule, and pack the pool swept by okc_match_to_map /
 * okc_match_to_map_uninitialised: rows[l] = o + 1 rows per visible landmark (landmarks ascending, 0 for invisible
 * ones), *nm rows in total, per row the 48-byte descriptor, owner (landmark index), e_W = C_WC_old * e_C.normalized(),
 * r_W = camera centre, filled (0: the slot was never written - the reference leaves such a row uninitialised, here it
 * is zero) and row_obs (index of the observation inside the landmark's list, -1). Output capacity: 3 rows per landmark.
 * An observation refers to view obs_view (one camera image of an older multiframe: pose, rows row0 .. of the shared
 * descriptor / back-projection pools) and keypoint obs_kp inside it. focal_sum = focalLengthU() + focalLengthV()
 * (:1361-1363). Host buffers. */
typedef struct okc_landmark_views {
  int n_views;
  const double* C_WC;   /* n_views x 9, row-major rotation of T_WC_old = T_WS_old * T_SC */
  const double* r_WC;   /* n_views x 3 */
  const int64_t* row0;  /* n_views: first pool row of the view (ascending) */
  const uint8_t* desc;  /* n_rows x 48 */
  const double* e_C;    /* n_rows x 3: Frame::backProjections_ (okc_back_project) */
  int64_t n_rows;
} okc_landmark_views;
int okc_prepare_landmarks(okc_context* ctx, const okc_landmark_views* views, int n_landmarks, const double* hp_W,
                          const double* quality, const uint8_t* visible, const int64_t* obs_offset,
                          const int32_t* obs_view, const int32_t* obs_kp, const double r_WC1[3], double focal_sum,
                          int loop_closure, uint8_t* is3d, int32_t* rows, int32_t* nm, uint8_t* pool, int32_t* owner,
                          double* e_W, double* r_W, uint8_t* filled, int32_t* row_obs);

/* Place recognition, front half: DBoW2::TemplatedVocabulary<FBrisk>::transform of every
 * descriptor of a frame (external/DBoW2/include/DBoW2/TemplatedVocabulary.h:1216-1257; distance = FBrisk::distance =
 * the 48-byte Hamming primitive, okvis_frontend/src/FBrisk.cpp:62-65), as TemplatedDatabase::query / add call it for
 * Frontend::getFilteredDBoWResult and the database fill (Frontend.cpp:605-671, 791-806).
 * okc_vocabulary_create: the n nodes of the vocabulary file (resources/small_voc.yml.gz: k = 9, L = 3, 819 nodes) in
 * FILE order - node_id (1..n, the root 0 is implicit), parent_id, 48-byte descriptor, weight, and word_of_node (the
 * word id of a leaf, -1 for inner nodes); the order of a node's children is the file order, as after load().
 * okc_vocabulary_transform: per feature the word id, its weight, and the node id at level `levels - levelsup`
 * (0 = root when that is <= 0), exactly the (word_id, weight, nid) of transform(). BowVector / FeatureVector
 * accumulation (std::map inserts, L1 normalisation) stays with the caller. Host buffers. */
typedef struct okc_vocabulary okc_vocabulary;
int okc_vocabulary_create(okc_context* ctx, int n, const int32_t* node_id, const int32_t* parent_id,
                          const uint8_t* descriptors, const double* weights, const int32_t* word_of_node, int depth,
                          okc_vocabulary** vocabulary);
int okc_vocabulary_destroy(okc_vocabulary* vocabulary);
int okc_vocabulary_transform(okc_vocabulary* vocabulary, const uint8_t* features, int n, int levels, int levelsup,
                             uint32_t* word_id, double* weight, uint32_t* node_id);

/* Data formats either side of the descriptor. Host buffers.
 * okc_descriptors_to_hex: the text form a saved map holds per keypoint, okvis_ceres/src/Component.cpp:434-437 -
 * 96 lower-case hex characters per 48-byte descriptor, two per byte, no separator and no terminator (hex[n*96]).
 * okc_descriptors_from_hex: the reader of the same file, Component.cpp:228-236 - every 2-character field goes
 * through `std::stringstream >> std::hex >> uint32_t` and the low byte is kept, so upper-case digits are accepted,
 * a field whose second character is no hex digit yields its first digit, a field without digits yields 0, and a
 * sign is honoured ("-1" -> 255); a field of white space only leaves the reference's byte uninitialised and is 0
 * here; hex[n*96] in, descriptors[n*48] out. The reference's std::out_of_range for a
 * string shorter than 96 characters is the adapter's length check. */
int okc_descriptors_to_hex(okc_context* ctx, const uint8_t* descriptors, int n, char* hex);
int okc_descriptors_from_hex(okc_context* ctx, const char* hex, int n, uint8_t* descriptors);
/* DBoW2 cluster centre, FBrisk::meanValue (okvis_frontend/src/FBrisk.cpp:25-60), for n_groups sets of descriptors at
 * once: group g = rows group_start[g] .. group_start[g+1]-1 of descriptors[.*48] (k-means step of vocabulary
 * training: one group per cluster). Bit b of mean[g*48 ..] is set when MORE than size/2 (integer division) rows
 * have it set; an empty group gives 48 zero bytes. */
int okc_descriptor_mean(okc_context* ctx, const uint8_t* descriptors, const int32_t* group_start, int n_groups,
                        uint8_t* mean);

/* ---- Frame pipeline: the per-frame sequence of Frontend::detectAndDescribe followed by Frontend::matchToMap
 * (ThreadedSlam.cpp:676-717 -> Frontend.cpp:204-262, 1335-1536) with every intermediate product left in HBM.
 * The queries of the map match are the camera handle's own descriptors/keypoints of the frame just described,
 * every keypoint is "unassociated" (a fresh frame has no landmark ids yet), and only the match result crosses
 * the bus. One stream per camera, as one std::thread per camera in the reference. ---- */

/* Same as okc_detect_describe_enqueue but the image is already in device memory (u8[H*W], stride == width). */
int okc_detect_describe_enqueue_device(okc_camera* cam, const void* image_device, const okc_detect_params* dp,
                                       const okc_describe_params* xp);

/* The descriptor pool of the landmarks to match (Frontend.cpp:1365-1411): rows map[nm*48] in landmark-major order
 * (std::map order), map_owner[m] = landmark slot of row m, ascending. Uploaded once, reused for every frame while
 * the keyframe window does not change. */
typedef struct okc_map_window okc_map_window;
int okc_map_window_create(okc_context* ctx, const uint8_t* map, const int32_t* map_owner, int nm, int num_landmarks,
                          okc_map_window** win);
int okc_map_window_destroy(okc_map_window* win);

/* Frontend::matchToMapByThread (Frontend.cpp:1744-1827) of the frame last enqueued on `cam` against `win`, on the
 * camera's stream. proj_xy: double[2*num_landmarks] landmark projections for this frame's pose, in host memory
 * (proj_on_device == 0; copied) or already in device memory (proj_on_device != 0; used in place until the finish).
 * num_slices =
 * num_matching_threads (<= 64): ctrs[] / repr_errors[] are accumulated per thread slice, in the reference's order, by a
 * kernel (match_stats) - the improvement events never leave the device. */
int okc_camera_match_to_map_enqueue(okc_camera* cam, const okc_map_window* win, const double* proj_xy,
                                    int proj_on_device, double thresh_sq, double hamming_threshold, int num_slices);
/* Waits for the camera's stream, then returns what okc_detect_describe_finish and okc_match_to_map return. */
int okc_camera_frame_finish(okc_camera* cam, const okc_map_window* win, okc_keypoint* keypoints, int capacity, int* n,
                            uint8_t* descriptors, double* best_dist, int32_t* best_owner, uint64_t* ctrs,
                            double* repr_errors);

/* ---- Batches: several images per launch chain. A batch has n_slots image slots; slot i carries an image of camera
 * cams[i] (a camera may own several slots: frames of a sequence; or the slots are the cameras of one multiframe,
 * okvis::MultiFrame, MultiFrame.hpp:307-312). One enqueue runs detect -> describe (-> matchToMap) for all slots with
 * ONE kernel launch per stage (grid.z = slot), which is how the 148 SMs are kept busy by 0.36 MB images; several
 * batches may be in flight at once (each owns a stream). Replaces the one-std::thread-per-camera fan-out of
 * ThreadedSlam.cpp:676-708 and the per-camera loop of Frontend::matchToMap (Frontend.cpp:1519-1536). ---- */
typedef struct okc_batch okc_batch;
int okc_batch_create(okc_context* ctx, okc_camera* const* cams, int n_slots, okc_batch** batch);
int okc_batch_destroy(okc_batch* batch);
int okc_batch_size(const okc_batch* batch);
/* The kernel chains of okc_batch_enqueue and okc_batch_match_landmarks_enqueue are replayed as CUDA graphs (captured
 * the second time a configuration is seen; one cudaGraphLaunch per chain instead of a launch per kernel): everything
 * that changes per frame reaches the kernels through device-resident tables. enable = 0 turns that off (plain
 * launches); the results are identical. Default: on. */
int okc_batch_set_graph(okc_batch* batch, int enable);
/* images[i]: u8[H*W]; host memory (pinned buffers are DMA'd in place, pageable ones are staged) or, with
 * images_on_device != 0, device memory. xp: n_slots describe parameter blocks (extraction direction per slot).
 * windows == NULL: no map match; else windows[i] (may be NULL per slot), proj_xy[i] = double[2*num_landmarks] in
 * host memory (copied) or, with proj_on_device != 0, device memory that is used in place and must stay valid and
 * unchanged until okc_batch_finish. thresh_sq[i] = squared pixel gate of slot i. Host images that are adjacent in
 * pinned memory (a ring of frames) are uploaded with one copy. */
int okc_batch_enqueue(okc_batch* batch, const uint8_t* const* images, int images_on_device,
                      const okc_detect_params* dp, const okc_describe_params* xp,
                      const okc_map_window* const* windows, const double* const* proj_xy, int proj_on_device,
                      const double* thresh_sq, double hamming_threshold, int num_slices);
/* Waits for the batch's stream and brings every slot's results to the host (two copies for the whole batch). */
int okc_batch_finish(okc_batch* batch);
/* After okc_batch_finish: what okc_detect_describe_finish / okc_match_to_map return, for one slot. */
int okc_batch_slot_frame(okc_batch* batch, int slot, okc_keypoint* keypoints, int capacity, int* n,
                         uint8_t* descriptors);
int okc_batch_slot_matches(okc_batch* batch, int slot, double* best_dist, int32_t* best_owner, uint64_t* ctrs,
                           double* repr_errors);
/* All slots at once into caller arrays with stride max_keypoints per slot: keypoints[n_slots*max_kp],
 * counts[n_slots], descriptors[n_slots*max_kp*48], and (NULL to skip) best_dist / best_owner [n_slots*max_kp],
 * ctrs / repr_errors [n_slots*num_slices]. */
int okc_batch_read_all(okc_batch* batch, okc_keypoint* keypoints, int32_t* counts, uint8_t* descriptors,
                       double* best_dist, int32_t* best_owner, uint64_t* ctrs, double* repr_errors);
void* okc_batch_stream(okc_batch* batch);

/* ---- Frame pipeline, the stages that follow matchToMapByThread in Frontend::dataAssociationAndInitialization
 * (Frontend.cpp:674-1143), device resident: a slot's descriptors, keypoints and back-projections are used where
 * describe left them, the landmark pool of the keyframe window stays in HBM, and per frame only the poses go in and
 * the match results come out. All calls below are enqueued on the batch's stream after okc_batch_enqueue of the same
 * frame; okc_batch_finish brings every result to the host (one copy per result block). ---- */

/* Geometry of a camera: enables Frame::computeBackProjections (okvis_cv/include/okvis/implementation/Frame.hpp:158-173)
 * as a stage after describe in every chain that runs on this camera (single-camera entry points and batches), and
 * gives the landmark stage its focal lengths. */
int okc_camera_set_geometry(okc_camera* cam, const okc_pinhole* model);

/* Pose of a slot's camera for the frame in flight: T_WC = T_WS * T_SC as rotation (row-major) + position, and its
 * inverse T_CW as rotation + translation (the reference forms both, Frontend.cpp:1356-1358; passing both keeps the
 * library free of a 3x3 inverse whose rounding would have to match Eigen's). */
typedef struct okc_slot_pose {
  double T_CW[12];
  double C_WC[9];
  double r_WC[3];
} okc_slot_pose;
int okc_batch_set_poses(okc_batch* batch, const okc_slot_pose* poses /* n_slots */);
/* After okc_batch_finish: e_C[n*3], valid[n] and the number of valid back-projections of a slot (n = its keypoints). */
int okc_batch_slot_back_projections(okc_batch* batch, int slot, double* e_C, uint8_t* valid, int* n_valid);

/* The landmarks of the keyframe window with their observations (what Frontend::matchToMap walks, Frontend.cpp:1335-1508),
 * uploaded once and kept in HBM: arguments as for okc_prepare_landmarks. okc_landmark_pool_update replaces the
 * landmark positions / qualities and the poses of the observing views after an optimisation (NULL = unchanged). */
typedef struct okc_landmark_pool okc_landmark_pool;
int okc_landmark_pool_create(okc_context* ctx, const okc_landmark_views* views, int n_landmarks, const double* hp_W,
                             const double* quality, const int64_t* obs_offset, const int32_t* obs_view,
                             const int32_t* obs_kp, okc_landmark_pool** pool);
int okc_landmark_pool_update(okc_landmark_pool* pool, const double* hp_W, const double* quality, const double* C_WC,
                             const double* r_WC);
int okc_landmark_pool_destroy(okc_landmark_pool* pool);

/* Frontend::matchToMap for every slot of the batch (Frontend.cpp:1335-1716 without the estimator calls): landmark
 * preparation for the slot's pose (projection, field-of-view test, is3d, viewpoint / scale pruning, best three
 * descriptors, pool packing: okc_project_landmarks + okc_prepare_landmarks as kernels on the batch's stream),
 * matchToMapByThread on the 3D landmarks and - with OKC_STAGE_UNINITIALISED - matchToMapByThreadUnitialised on the
 * others, where a keypoint takes part when its back-projection is valid and matchToMapByThread left it without a
 * landmark (a fresh frame has no other landmark ids). The reprojection gate is 3 + 0.06 f (3 + 0.34 f with
 * OKC_STAGE_NO_IMU), f = 0.5 (fu + fv) of the slot's camera geometry (Frontend.cpp:1340-1342, 1760-1764).
 * Needs okc_camera_set_geometry on every slot's camera and okc_batch_set_poses. */
#define OKC_STAGE_UNINITIALISED 1
#define OKC_STAGE_NO_IMU 2
#define OKC_STAGE_LOOP_CLOSURE 4
int okc_batch_match_landmarks_enqueue(okc_batch* batch, const okc_landmark_pool* pool, double hamming_threshold,
                                      int num_slices, int flags);
/* After okc_batch_finish. okc_batch_slot_matches returns the matchToMapByThread part (best_owner = landmark index).
 * okc_batch_slot_uninitialised: what okc_match_to_map_uninitialised returns, for the slot's n keypoints.
 * okc_batch_slot_landmarks: the prepared pool of a slot (copied from the device on demand; any pointer may be NULL):
 * proj_xy[L*2], visible[L], is3d[L], rows[L], *nm and per row owner[nm], pool[nm*48]. */
int okc_batch_slot_uninitialised(okc_batch* batch, int slot, double* best_dist, int32_t* best_owner, double* hps_W,
                                 uint8_t* hps_written, uint64_t* ctrs);
int okc_batch_slot_landmarks(okc_batch* batch, int slot, double* proj_xy, uint8_t* visible, uint8_t* is3d,
                             int32_t* rows, int32_t* nm, int32_t* owner, uint8_t* pool);

/* Frontend::matchStereo (Frontend.cpp:2259-2316) between slots of the batch (the cameras of one multiframe):
 * pair p matches the keypoints of slot0[p] (k0) against those of slot1[p] (k1). Focal length f = 0.5 (fu + fv) of the
 * cameras' geometry. After okc_batch_finish, okc_batch_pair_stereo returns what okc_match_stereo returns. */
int okc_batch_match_stereo_enqueue(okc_batch* batch, int n_pairs, const int* slot0, const int* slot1,
                                   double hamming_threshold);
int okc_batch_pair_stereo(okc_batch* batch, int pair, double* dist, int32_t* k1_match, double* hp_W,
                          uint8_t* initialisable);

/* Camera-sharded multiframes (one GPU per camera subset): matchStereo needs both cameras of an overlapping pair, so
 * the ranks exchange their views. okc_batch_pack_views copies the views of `n` slots (keypoint count, keypoints,
 * descriptors, back-projections, validity flags) into consecutive records of okc_packed_view_bytes(max_keypoints)
 * bytes in caller-provided DEVICE memory, on the batch's stream - the send buffer of one NCCL all-gather.
 * okc_batch_match_stereo_packed_enqueue then matches slot0[p] (k0) against record packed_index[p] of the gathered
 * buffer (k1), whose camera has pose (C_WC, r_WC) = poses1[p] and focal length f1[p]. Results: okc_batch_pair_stereo. */
size_t okc_packed_view_bytes(int max_keypoints);
int okc_batch_pack_views(okc_batch* batch, int n, const int* slots, void* packed_device);
int okc_batch_match_stereo_packed_enqueue(okc_batch* batch, int n_pairs, const int* slot0, const void* packed_device,
                                          const int* packed_index, const okc_slot_pose* poses1, const double* f1,
                                          double hamming_threshold);

/* An older multiframe kept on the device for Frontend::matchMotionStereo (Frontend.cpp:1968-2150): okc_batch_retain
 * copies the descriptors, keypoints and back-projections of n_views slots (device to device, on the batch's stream)
 * after the frame was described; okc_keyframe_set_view gives a view its current pose (after an optimisation), focal
 * length and - optionally - the eligibility flags of its keypoints (use0: not an initialised landmark, not observed;
 * estimator state, Frontend.cpp:2056-2084; NULL = all eligible). */
typedef struct okc_keyframe okc_keyframe;
int okc_keyframe_create(okc_context* ctx, int n_views, int max_keypoints, okc_keyframe** keyframe);
int okc_keyframe_destroy(okc_keyframe* keyframe);
int okc_batch_retain(okc_batch* batch, int n_views, const int* slots, okc_keyframe* keyframe);
int okc_keyframe_set_view(okc_keyframe* keyframe, int view, const double C_WC[9], const double r_WC[3], double f,
                          const uint8_t* use0, int n_use0);
/* Job j: the keypoints of view views[j] of keyframes[j] (k0) against the keypoints of slot slots[j] that have no
 * landmark after the landmark stage (only_unmatched != 0; the reference packs them, Frontend.cpp:2034-2047) or against
 * all of them. k1_match is the keypoint index inside the slot (the reference's k1s[k1]). One call per frame with every
 * job (okc_batch_match_stereo_enqueue likewise); a keyframe can be matched once the frame it was retained from is
 * finished. After okc_batch_finish,
 * okc_batch_job_motion returns what okc_match_motion_stereo returns, for the view's keypoints. */
int okc_batch_match_motion_enqueue(okc_batch* batch, int n_jobs, const okc_keyframe* const* keyframes, const int* views,
                                   const int* slots, double hamming_threshold, int only_unmatched);
int okc_batch_job_motion(okc_batch* batch, int job, int* n0, double* dist, int32_t* k1_match, double* hp_W,
                         uint8_t* initialisable, double* quality);
/* All stage results of a finished frame at once into caller arrays (a group whose first pointer is NULL is skipped):
 * uninitialised-landmark matches per slot with stride max_keypoints (u_ctrs: num_slices per slot), stereo pairs with
 * stride max_keypoints, motion jobs with stride = the largest max_keypoints among the batch and the keyframes
 * (m_n0[j] = keypoints of job j's keyframe view). */
int okc_batch_read_stages(okc_batch* batch, double* u_dist, int32_t* u_owner, double* u_hps, uint8_t* u_written,
                          uint64_t* u_ctrs, double* s_dist, int32_t* s_k1, double* s_hp, uint8_t* s_init, int32_t* m_n0,
                          double* m_dist, int32_t* m_k1, double* m_hp, uint8_t* m_init, double* m_quality);

/* The same results WITHOUT a copy: okc_batch_finish leaves everything in pinned host memory owned by the batch, and
 * this returns where (the way the reference's callers read keypoints_ / descriptors_ / the match vectors in place
 * out of okvis::Frame). Contents are valid from okc_batch_finish until the next enqueue on this batch; the pointers
 * until the stage set-up grows (more pairs / jobs / landmarks) or the batch is destroyed. Rows beyond a count are
 * unspecified. A group whose base pointer is NULL did not run for this frame. */
typedef struct okc_match_stats { uint64_t count; double sum; } okc_match_stats; /* ctrs[t], reprErrors[t] * ctrs[t] */
typedef struct okc_job_view {      /* result blocks of matchStereo pairs or matchMotionStereo jobs */
  const uint8_t* base;             /* block j at base + j * stride */
  size_t stride;
  size_t off_dist, off_k1, off_hp, off_initialisable, off_quality; /* double[rows], int32[rows], double[rows][4], u8[rows], double[rows] (motion only) */
  int n, rows;                     /* blocks, row capacity of a block */
} okc_job_view;
typedef struct okc_batch_view {
  int n_slots, max_keypoints, num_slices;
  const int32_t* counters;         /* [n_slots][8]: [2] = keypoints after describe (live rows of every per-keypoint array) */
  const uint8_t* slots;            /* slot i at slots + i * slot_stride */
  size_t slot_stride;
  size_t off_keypoints, off_descriptors;      /* okc_keypoint[max_keypoints], u8[max_keypoints][48] */
  size_t off_best_dist, off_best_owner;       /* matchToMapByThread: double[max_keypoints], int32[max_keypoints] */
  size_t off_stats;                           /* okc_match_stats[num_slices] */
  int map_match;                              /* 0: no map match ran (the three offsets above are meaningless) */
  const uint8_t* uninit;           /* matchToMapByThreadUnitialised: slot i at uninit + i * uninit_stride, or NULL */
  size_t uninit_stride;
  size_t u_off_dist, u_off_owner, u_off_hps, u_off_written; /* double[], int32[], double[][4], u8[] per keypoint */
  const uint64_t* uninit_ctrs;     /* [n_slots][uninit_ctrs_stride], the first num_slices of a row are used */
  int uninit_ctrs_stride;
  okc_job_view stereo;             /* rows of pair p: counters[8 * slot0 + 2] of its first slot */
  okc_job_view motion;             /* rows of job j: motion_n0[j] */
  const int32_t* motion_n0;
} okc_batch_view;
int okc_batch_view_results(okc_batch* batch, okc_batch_view* view);
/* Measurement hook: the bytes the copies of the last frame (enqueue .. finish) moved host->device and device->host. */
int okc_batch_transfer_bytes(const okc_batch* batch, uint64_t* h2d, uint64_t* d2h);

int okc_debug_batch_kernel_timing(okc_batch* batch, int enable);
int okc_debug_batch_kernel_times(okc_batch* batch, char* names, int names_cap, float* ms, int cap, int* n);

/* Measurement hooks: the camera's cudaStream_t (to bracket enqueued work with CUDA events from a harness), and
 * per-kernel device times of the last frame enqueued while timing was enabled (CUDA events around every launch on
 * the camera's stream). names: '\n'-separated kernel labels. */
void* okc_camera_stream(okc_camera* cam);
int okc_debug_kernel_timing(okc_camera* cam, int enable);
int okc_debug_kernel_times(okc_camera* cam, char* names, int names_cap, float* ms, int cap, int* n);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* OKVIS_CUDA_H_ */
