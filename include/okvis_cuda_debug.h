/* okvis_cuda - test hooks and measurement aids. NOT part of the drop-in boundary: nothing in OKVIS2-X calls these.
 *
 * The okc_debug_set_* switches select the A/B baselines the parity tests compare the default kernels with (each
 * baseline follows the reference loop more literally) or force a launch shape. They are process-wide and read at
 * launch time: set them only while no other okvis_cuda call is in flight. Production callers never touch them.
 */
#ifndef OKVIS_CUDA_DEBUG_H_
#define OKVIS_CUDA_DEBUG_H_

#include "okvis_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

/* Uniformity selection (uniformity-enforcement-inl.h:57-207): 0 = wave kernel (default), 1 = sequential-visit kernel. */
int okc_debug_set_select_wave(int variant);
/* Image rows one Harris warp walks: 0 = chosen per launch by the cost model. */
int okc_debug_set_harris_rows(int rows);
/* Plain best match: 0 = lower-bound sweep kernel (default), 12 = plain 12-word Hamming sweep. */
int okc_debug_set_match_prefix(int words);
/* Frame-pipeline map match: 1 = through spatial bins of the projections (default), 0 = scan of every map row. */
int okc_debug_set_match_binned(int on);

/* Counter `which` (0 candidates, 1 kept, 2 described, 4 match events, 5 selection rounds) of a slot, as read back by
 * the last okc_batch_finish. */
int okc_debug_batch_counter(okc_batch* batch, int slot, int which, int* value);

/* How many cudaMalloc / cudaMallocHost calls (allocations) and cudaFree / cudaFreeHost calls (releases) the library
 * has made in this process so far; every such call of the library is counted. The per-frame entry points of a batch
 * (okc_batch_enqueue*, okc_batch_set_poses, okc_batch_match_*_enqueue, okc_batch_finish, okc_batch_view_results)
 * leave both numbers unchanged once the batch has seen its largest inputs: tests/test_gpu_pipeline.py checks that. */
int okc_debug_alloc_counts(uint64_t* allocations, uint64_t* releases);

/* The reference's std::sort order of `n` PointWithScore-like scores (raster order in), for the tie-order tests. */
int okc_debug_std_sort(okc_camera* cam, const int32_t* scores, int n, uint32_t* order);

/* The host-built pattern tables (no device work): points float[1024*66*3], long_pairs int32[856*4] = (i, j, wdx, wdy),
 * short_pairs u8[384*2], lut float[31*31]; any pointer may be NULL. CPU-only tests pin them against the reference's
 * constructor (brisk/src/brisk-descriptor-extractor.cc:189-340). */
int okc_debug_host_tables(float* points, int* basicscale, int* border, int32_t* long_pairs, uint8_t* short_pairs,
                          float* lut);

/* Device pointer / size of a descriptor set (harnesses that time kernels with inputs resident in HBM). */
void* okc_descriptor_set_device_ptr(const okc_descriptor_set* set);
int64_t okc_descriptor_set_size(const okc_descriptor_set* set);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* OKVIS_CUDA_DEBUG_H_ */
