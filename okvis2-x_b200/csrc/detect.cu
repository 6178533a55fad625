// Keypoint detection kernels.
//
//  harris_nms_kernel   = brisk::HarrisScoresSSE (brisk/src/harris-scores.cc:65-301) fused with
//                        HarrisScoreCalculator::Get2dMaxima (brisk/src/harris-score-calculator.cc:70-119):
//                        one pass over the u8 image, no int16 planes and no score map in HBM; writes a bitmap of
//                        the NMS survivors and their scores, compacted per (row, strip) segment.
//  order_candidates_kernel = the std::sort at the top of EnforceKeyPointUniformity, tie order included
//                        (libstdc++ introsort reproduced exactly; see the comment at the kernel).
//  select_keypoints_wave_kernel = the greedy loop of EnforceKeyPointUniformity
//                        (brisk/include/brisk/internal/uniformity-enforcement-inl.h:57-207) in its dependency
//                        form (default); select_keypoints_kernel = the same loop one candidate at a time with the
//                        occupancy image in shared memory (A/B baseline, and for more keypoints per image than the
//                        dependency form's lists hold).
//  subpixel_kernel     = the Subpixel2D tail of DetectScaleSpaceMaxima
//                        (brisk/include/brisk/internal/scale-space-layer-inl.h:394-420, 568-701); the 3x3 scores
//                        around a kept point are recomputed from the pixels (harris_ring_scores).
#include "common.cuh"
#include <cstdio>
#include <cstdlib>
#include <type_traits>

namespace okc {

// ----------------------------------------------------------------------------------------------
// Harris score + 8-neighbour NMS
// ----------------------------------------------------------------------------------------------
// Register-marching stencil: a warp owns a strip of 128 columns (4 consecutive columns per lane) and walks down
// rowsPerWarp output rows; every intermediate (Scharr partial sums, gradient products, horizontally smoothed
// products, scores, row maxima) lives in registers, neighbours across lanes come from warp shuffles, and there is
// no shared memory and no block-level synchronisation. Lanes 0 and 31 only feed their neighbours (halo), so a
// warp emits 120 columns. Per pixel this is ~60 integer operations against ~320 in a shared-memory tile
// formulation (index arithmetic, byte accesses, four barriers) - the kernel is issue-bound, so that is the time.
//
// Arithmetic (brisk/src/harris-scores.cc:124-296): with h = I[x-1] - I[x+1] and s = 3 I[x-1] + 10 I[x] + 3 I[x+1]
// per row, dx = (3 h[y-1] + 10 h[y] + 3 h[y+1]) << 3, dy = (s[y-1] - s[y+1]) << 3 (both fit int16 as in the
// reference), products (a*b) >> 16 (= _mm_mulhi_epi16), 3x3 binomial smoothing as one integer sum >> 4, score =
// dxdx*dydy - dxdy^2 - (((dxdx + dydy) >> 1)^2) / 5. Scores are defined on rows 2..H-3, cols 2..W-3 and 0 on the
// frame (Mat::zeros, :73).
//
// NMS survivors (harris-score-calculator.cc:76-117: >= threshold, no neighbour strictly greater) are published as a
// bitmap with kStripWords words per (row, strip): bit b of word j of strip s is column 120 s - 4 + 32 j + b.
// Row-major, strip-major, bit-ascending order = raster order, which order_candidates_kernel relies on.
// output rows per warp: chosen per launch (harris_rows_per_warp), see launch_harris_nms
constexpr int kHarrisWarps = 1;      // warps per block (independent: no shared memory, no barriers)
constexpr int kStripCols = 120;      // output columns per warp
constexpr int kStripWords = 4;

__host__ __device__ inline int harris_strips(int W) { return (W + kStripCols - 1) / kStripCols; }

//
// Score map: the only scores anything reads back are those of the NMS survivors (order_candidates_kernel builds its
// sort keys from them); the 3x3 ring of the sub-pixel fit is recomputed from the pixels for the few hundred KEPT
// points (harris_ring_scores). So by default (DENSE = false) only the survivors' scores are stored, compacted per
// (row, strip) segment: segment (q, strip) holds its survivors' scores in column order at
// scores[(q * strips + strip) * kStripCols ...] (harris_score_ints() ints per image) - 4 bytes per candidate instead
// of 4 bytes per pixel, which was 2/3 of the kernel's DRAM traffic. DENSE = true writes the
// whole map: scale-space layers (scale_nms_kernel interpolates in the neighbouring layers' maps) and okc_debug_scores.
template <bool ALIGNED, bool DENSE>
__global__ void __launch_bounds__(32 * kHarrisWarps) harris_nms_kernel(const Slot* __restrict__ slots, int W, int H,
                                                         int absThreshold, int rowsPerWarp) {
  const Slot& slot = slots[blockIdx.z];
  const uint8_t* __restrict__ img = slot.image;
  int32_t* __restrict__ scores = slot.scores;
  uint32_t* __restrict__ bitmap = slot.bitmap;
  const int lane = threadIdx.x & 31;
  const int strip = blockIdx.x;
  const int q0 = (blockIdx.y * kHarrisWarps + (threadIdx.x >> 5)) * rowsPerWarp;
  if (q0 >= H) return;
  const int q1 = min(H, q0 + rowsPerWarp);
  const int x = strip * kStripCols - 4 + 4 * lane;  // first of this lane's 4 columns
  const int wordsPerRow = harris_strips(W) * kStripWords;
  const bool emit = lane >= 1 && lane <= 30;

  // Row histories: slot [P] receives the row computed in a step, [(P+1)%3] holds the row before the previous one ("A"),
  // [(P+2)%3] the previous one ("B"). P advances with the row, the loop is unrolled by 3 and the slots are fixed
  // registers: no register-to-register rotation (it was 110 of the 382 instructions of a step).
  int hH[3][4] = {}, sH[3][4] = {}, xxH[3][4] = {}, xyH[3][4] = {}, yyH[3][4] = {}, mH[3][4] = {}, scH[3][4] = {},
      lrH[3][4] = {};

  // pixels of row r (clamped; rows/columns outside the image never reach a valid score)
  auto load_row = [&](int r) -> unsigned {
    const int rr = r < 0 ? 0 : (r >= H ? H - 1 : r);
    unsigned w = 0;
    if (ALIGNED) {
      if (x >= 0 && x < W) w = __ldg(reinterpret_cast<const unsigned*>(img + (size_t)rr * W + x));
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (x + c >= 0 && x + c < W) w |= (unsigned)__ldg(img + (size_t)rr * W + x + c) << (8 * c);
    }
    return w;
  };
  // columns 2 .. W-3 carry a score (lane constants): all ones / zero per column, and as a 4-bit mask
  int colKeep[4];
  unsigned colBits = 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    colKeep[c] = (x + c >= 2 && x + c <= W - 3) ? -1 : 0;
    colBits |= (colKeep[c] & 1u) << c;
  }
  constexpr int kAhead = 5;   // rows in flight
  unsigned wq[kAhead];
#pragma unroll
  for (int i = 0; i < kAhead; ++i) wq[i] = load_row(q0 - 3 + i);
  auto step = [&](auto Pc, int r) {
    constexpr int P = decltype(Pc)::value, A = (P + 1) % 3, B = (P + 2) % 3;
    // ---- 1. pixels of row r; rows r+1 .. r+kAhead-1 are already in flight (the pixels come from DRAM: a step is
    // ~300 instructions, three steps of the 4 warps of a scheduler cover its latency) ----
    const unsigned w = wq[0];
#pragma unroll
    for (int i = 0; i + 1 < kAhead; ++i) wq[i] = wq[i + 1];
    wq[kAhead - 1] = load_row(r + kAhead);
    const unsigned wl = __shfl_up_sync(0xffffffffu, w, 1), wr = __shfl_down_sync(0xffffffffu, w, 1);
    int px[6];
    px[0] = (int)(wl >> 24);
    px[1] = (int)(w & 0xff);
    px[2] = (int)((w >> 8) & 0xff);
    px[3] = (int)((w >> 16) & 0xff);
    px[4] = (int)(w >> 24);
    px[5] = (int)(wr & 0xff);
    // ---- 2. gradients and products of row r-1, horizontally smoothed ----
    int pxx[4], pxy[4], pyy[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      // Scharr with the reference's "<< 3" folded into the weights (|dx|, |dy| <= 32640: no int16 wrap to mirror);
      // the smoothed-row history sH holds 8 * (3, 10, 3)-sums
      const int h2 = px[c] - px[c + 2];
      const int s2 = 24 * (px[c] + px[c + 2]) + 80 * px[c + 1];
      const int dx = 24 * (hH[A][c] + h2) + 80 * hH[B][c];
      const int dy = sH[A][c] - s2;
      hH[P][c] = h2;
      sH[P][c] = s2;
      pxx[c] = (dx * dx) >> 16;
      pxy[c] = (dx * dy) >> 16;
      pyy[c] = (dy * dy) >> 16;
    }
    int xx2[4], xy2[4], yy2[4];
    {
      const int lxx = __shfl_up_sync(0xffffffffu, pxx[3], 1), rxx = __shfl_down_sync(0xffffffffu, pxx[0], 1);
      const int lxy = __shfl_up_sync(0xffffffffu, pxy[3], 1), rxy = __shfl_down_sync(0xffffffffu, pxy[0], 1);
      const int lyy = __shfl_up_sync(0xffffffffu, pyy[3], 1), ryy = __shfl_down_sync(0xffffffffu, pyy[0], 1);
      xx2[0] = lxx + 2 * pxx[0] + pxx[1];
      xx2[1] = pxx[0] + 2 * pxx[1] + pxx[2];
      xx2[2] = pxx[1] + 2 * pxx[2] + pxx[3];
      xx2[3] = pxx[2] + 2 * pxx[3] + rxx;
      xy2[0] = lxy + 2 * pxy[0] + pxy[1];
      xy2[1] = pxy[0] + 2 * pxy[1] + pxy[2];
      xy2[2] = pxy[1] + 2 * pxy[2] + pxy[3];
      xy2[3] = pxy[2] + 2 * pxy[3] + rxy;
      yy2[0] = lyy + 2 * pyy[0] + pyy[1];
      yy2[1] = pyy[0] + 2 * pyy[1] + pyy[2];
      yy2[2] = pyy[1] + 2 * pyy[2] + pyy[3];
      yy2[3] = pyy[2] + 2 * pyy[3] + ryy;
    }
    // ---- 3. score of row p = r-2 ----
    const int p = r - 2;
    const int rowKeep = (p >= 2 && p <= H - 3) ? -1 : 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int a = (xxH[A][c] + 2 * xxH[B][c] + xx2[c]) >> 4;
      const int b = (yyH[A][c] + 2 * yyH[B][c] + yy2[c]) >> 4;
      const int d = (xyH[A][c] + 2 * xyH[B][c] + xy2[c]) >> 4;
      const int t = (a + b) >> 1;
      const int v = a * b - d * d - (int)((unsigned)(t * t) / 5u);
      xxH[P][c] = xx2[c];
      xyH[P][c] = xy2[c];
      yyH[P][c] = yy2[c];
      scH[P][c] = v & rowKeep & colKeep[c];
    }
    if (DENSE && emit && p >= q0 && p < q1 && x < W) {
      int32_t* dst = scores + (size_t)p * W + x;
      if (ALIGNED) {
        *reinterpret_cast<int4*>(dst) = make_int4(scH[P][0], scH[P][1], scH[P][2], scH[P][3]);
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (x + c < W) dst[c] = scH[P][c];
      }
    }
    const int ls = __shfl_up_sync(0xffffffffu, scH[P][3], 1), rs = __shfl_down_sync(0xffffffffu, scH[P][0], 1);
    lrH[P][0] = max(ls, scH[P][1]);
    lrH[P][1] = max(scH[P][0], scH[P][2]);
    lrH[P][2] = max(scH[P][1], scH[P][3]);
    lrH[P][3] = max(scH[P][2], rs);
    int mTop[4];   // row r-4, read before slot P of the maxima history is overwritten
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      mTop[c] = mH[A][c];
      mH[P][c] = max(lrH[P][c], scH[P][c]);
    }
    // ---- 4. NMS of row q = r-3: top = row r-4, middle = row r-3 (slot B), bottom = row r-2 (slot P) ----
    const int q = r - 3;
    if (q >= q0 && q < q1) {   // warp-uniform
      // a row segment without a single score at the threshold has no maxima: one vote instead of the comparisons
      // and the three shuffles that assemble the bitmap words
      const int rowMax = max(max(scH[B][0], scH[B][1]), max(scH[B][2], scH[B][3]));
      unsigned v = 0;
      if (__any_sync(0xffffffffu, rowMax >= absThreshold)) {
        unsigned nib = 0;
        if (emit && q >= 2 && q <= H - 3) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            // >= the threshold and >= all 8 neighbours, as one comparison against the largest of them
            const bool isMax = scH[B][c] >= max(max(mTop[c], mH[P][c]), max(lrH[B][c], absThreshold));
            nib |= (isMax ? 1u : 0u) << c;
          }
          nib &= colBits;
        }
        v = nib << (4 * (lane & 7));
        v |= __shfl_xor_sync(0xffffffffu, v, 1);
        v |= __shfl_xor_sync(0xffffffffu, v, 2);
        v |= __shfl_xor_sync(0xffffffffu, v, 4);
        if (!DENSE) {
          // the survivors' scores, compacted per (row, strip) segment in column order: a segment's scores share
          // 32-byte sectors (scattered 4-byte stores into a dense map cost a sector each, in the store and again in
          // order_candidates_kernel's load)
          const int cw = __popc(v);   // survivors of this lane's word (the same in its 8 lanes)
          const int c0 = __shfl_sync(0xffffffffu, cw, 0), c1 = __shfl_sync(0xffffffffu, cw, 8),
                    c2 = __shfl_sync(0xffffffffu, cw, 16);
          if (nib) {
            const int g = lane >> 3;
            int rank = (g > 0 ? c0 : 0) + (g > 1 ? c1 : 0) + (g > 2 ? c2 : 0) + __popc(v & ((1u << (4 * (lane & 7))) - 1u));
            int32_t* seg = scores + ((size_t)q * gridDim.x + strip) * kStripCols;
#pragma unroll
            for (int c = 0; c < 4; ++c)
              if (nib & (1u << c)) seg[rank++] = scH[B][c];
          }
        }
      }
      if ((lane & 7) == 0) bitmap[(size_t)q * wordsPerRow + strip * kStripWords + (lane >> 3)] = v;
    }
  };
  // rows q0-3 .. q1+2, three per trip (an extra step or two past the end only see clamped rows and emit nothing)
  for (int r = q0 - 3; r <= q1 + 2; r += 3) {
    step(std::integral_constant<int, 0>{}, r);
    step(std::integral_constant<int, 1>{}, r + 1);
    step(std::integral_constant<int, 2>{}, r + 2);
  }
}

// Rows per warp: a warp computes rows + 6 halo rows (rounded up to the loop's 3 rows per trip), so tall chunks waste
// less, but the grid must still fill the GPU in whole waves (16 one-warp CTAs of 128 registers per SM) and a
// small batch wants short chunks for latency. Cost model: trips per warp x waves of CTAs.
static int harris_rows_per_warp(int W, int H, int n) {
  if (g_knobs.harris_rows > 0) return g_knobs.harris_rows;   // tests: forced rows per warp
  static int slotsOf[64] = {};   // resident one-warp CTAs per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (slotsOf[dev & 63] == 0) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    slotsOf[dev & 63] = sms * (16 / kHarrisWarps);
  }
  const int slots = slotsOf[dev & 63];
  const int candidates[] = {8, 12, 16, 24, 32, 48, 64, 96, 120, 160, 240};
  int best = 32;
  long long bestCost = -1;
  for (int rows : candidates) {
    const int chunks = (H + rows - 1) / rows;
    const long long ctas = (long long)harris_strips(W) * ((chunks + kHarrisWarps - 1) / kHarrisWarps) * n;
    const long long waves = (ctas + slots - 1) / slots;
    const long long trips = (rows + 6 + 2) / 3;
    const long long cost = trips * waves;
    if (bestCost < 0 || cost < bestCost) {
      bestCost = cost;
      best = rows;
    }
  }
  return best;
}

void launch_harris_nms(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s, bool aligned) {
  const int rows = harris_rows_per_warp(a.width, a.height, n);
  const int chunks = (a.height + rows - 1) / rows;
  dim3 grid(harris_strips(a.width), (chunks + kHarrisWarps - 1) / kHarrisWarps, n);
  auto k = a.dense_scores ? (aligned ? harris_nms_kernel<true, true> : harris_nms_kernel<false, true>)
                          : (aligned ? harris_nms_kernel<true, false> : harris_nms_kernel<false, false>);
  k<<<grid, 32 * kHarrisWarps, 0, s>>>(slots, a.width, a.height, a.abs_threshold, rows);
}

// ints of a slot's score buffer: the dense map (W * H) or the per-segment compact form, whichever is larger
size_t harris_score_ints(int W, int H) {
  const size_t seg = (size_t)H * harris_strips(W) * kStripCols, px = (size_t)W * H;
  return seg > px ? seg : px;
}

size_t harris_bitmap_words(int W, int H) { return (size_t)H * harris_strips(W) * kStripWords; }

// position of pixel (x, y) in the bitmap (host helper for tests / debug entry points)
void harris_bitmap_pos(int W, int x, int y, size_t* word, int* bit) {
  const int strip = x / kStripCols, off = x - (strip * kStripCols - 4);
  *word = (size_t)y * harris_strips(W) * kStripWords + (size_t)strip * kStripWords + (off >> 5);
  *bit = off & 31;
}

// ----------------------------------------------------------------------------------------------
// order_candidates: raster-ordered candidate list + the reference's std::sort order
// ----------------------------------------------------------------------------------------------
// EnforceKeyPointUniformity starts with std::sort(points.begin(), points.end()) over the raster-ordered
// Get2dMaxima output with PointWithScore::operator< = "score > other.score"
// (uniformity-enforcement-inl.h:67, score-calculator.h:94-96). Scores tie often (they are quantised by the >>16
// and >>4 shifts), std::sort is not stable, and the order of equal scores decides both which points the greedy
// loop keeps and the index order of the output keypoints. So the device reproduces libstdc++'s introsort exactly:
//   * __introsort_loop = a tree of Hoare partitions (median-of-3 pivot moved to the front, unguarded partition,
//     recursion while a range has > 16 elements, heapsort when the depth limit 2*floor(log2 n) is exhausted).
//     Ranges at the same depth are disjoint, so they are processed level by level, one warp per range; the
//     sequential two-pointer partition is replaced by its closed form: the i-th "left stop" (score <= pivot,
//     scanning up) is swapped with the i-th "right stop" (score >= pivot, scanning down) while L_i < R_i.
//   * __final_insertion_sort is a stable insertion sort, i.e. the final order = a stable sort by score of the
//     array the partition phase leaves behind. That array is a chain of segments between consecutive cuts (at most
//     16 elements each, or heap-sorted) in which no element of a later segment belongs before an earlier one, so
//     every segment is sorted on its own, by ranking its elements in parallel (a global bitonic sort of 64-bit keys
//     was 40 % of the kernel).
// tools/introsort_model.py checks this formulation against the real std::sort (oracle/_ref) on tie-heavy and
// adversarial (heapsort-triggering) inputs.
constexpr int kOrdThreads = 1024;
constexpr int kOrdSmemCap = 11264;                        // candidates handled entirely in shared memory (188 KB)
constexpr int kOrdQueueSmem = 704;                        // >= kOrdSmemCap / 17 ranges per level
constexpr int kOrdQueueGlobal = kCandCap / 16;            // >= kCandCap / 17

struct OrdArrays {
  int* sc;          // score
  uint32_t* ra;     // raster index y * W + x
  int* lp;          // left stops of the range being partitioned (positions), stored inside the range
  int* rp;          // right stops
  int2* q[2];       // range queues of the current / next level
};

__device__ __forceinline__ void ord_swap(const OrdArrays& A, int i, int j) {
  const int s = A.sc[i];
  A.sc[i] = A.sc[j];
  A.sc[j] = s;
  const uint32_t r = A.ra[i];
  A.ra[i] = A.ra[j];
  A.ra[j] = r;
}

// std::__partial_sort(first, last, last) = make_heap + sort_heap with comp(x, y) = x.score > y.score
// (bits/stl_heap.h: __adjust_heap / __push_heap). Sequential; only reached on adversarial inputs.
__device__ __noinline__ void ord_heap_sort(const OrdArrays& A, int f, int l) {
  const int n = l - f;
  if (n < 2) return;
  auto adjust = [&](int hole, int length, int vs, uint32_t vr) {
    const int top = hole;
    int child = hole;
    while (child < (length - 1) / 2) {
      child = 2 * (child + 1);
      if (A.sc[f + child] > A.sc[f + child - 1]) --child;
      A.sc[f + hole] = A.sc[f + child];
      A.ra[f + hole] = A.ra[f + child];
      hole = child;
    }
    if ((length & 1) == 0 && child == (length - 2) / 2) {
      child = 2 * (child + 1);
      A.sc[f + hole] = A.sc[f + child - 1];
      A.ra[f + hole] = A.ra[f + child - 1];
      hole = child - 1;
    }
    int parent = (hole - 1) / 2;
    while (hole > top && A.sc[f + parent] > vs) {
      A.sc[f + hole] = A.sc[f + parent];
      A.ra[f + hole] = A.ra[f + parent];
      hole = parent;
      parent = (hole - 1) / 2;
    }
    A.sc[f + hole] = vs;
    A.ra[f + hole] = vr;
  };
  for (int parent = (n - 2) / 2;; --parent) {
    adjust(parent, n, A.sc[f + parent], A.ra[f + parent]);
    if (parent == 0) break;
  }
  for (int last = n - 1; last > 0; --last) {
    const int vs = A.sc[f + last];
    const uint32_t vr = A.ra[f + last];
    A.sc[f + last] = A.sc[f];
    A.ra[f + last] = A.ra[f];
    adjust(0, last, vs, vr);
  }
}

// std::__unguarded_partition_pivot on [f, l) by one warp; returns the cut.
__device__ __forceinline__ int ord_warp_partition(const OrdArrays& A, int f, int l, int lane) {
  if (lane == 0) {  // __move_median_to_first(first, first + 1, mid, last - 1)
    const int a = f + 1, b = f + (l - f) / 2, c = l - 1;
    const int sa = A.sc[a], sb = A.sc[b], s_c = A.sc[c];
    int pick;
    if (sa > sb) pick = (sb > s_c) ? b : ((sa > s_c) ? c : a);
    else pick = (sa > s_c) ? a : ((sb > s_c) ? c : b);
    ord_swap(A, f, pick);
  }
  __syncwarp();
  const int pv = A.sc[f];
  const int len = l - f;
  const unsigned lt = (1u << lane) - 1u;
  int nL = 0, nR = 0;
  for (int base = 0; base < len; base += 32) {
    const int i = base + lane;
    const int pL = f + 1 + i, pR = l - 1 - i;
    const bool isL = pL < l && !(A.sc[pL] > pv);   // "while (comp(*first, pivot)) ++first" stops here
    const bool isR = i < len && !(pv > A.sc[pR]);  // "while (comp(pivot, *last)) --last" stops here (f = sentinel)
    const unsigned bL = __ballot_sync(0xffffffffu, isL), bR = __ballot_sync(0xffffffffu, isR);
    if (isL) A.lp[f + 1 + nL + __popc(bL & lt)] = pL;
    if (isR) A.rp[f + nR + __popc(bR & lt)] = pR;
    nL += __popc(bL);
    nR += __popc(bR);
  }
  __syncwarp();
  const int m = nL < nR ? nL : nR;
  int t = 0;  // number of swaps: leading pairs with L_i < R_i
  for (int base = 0; base < m; base += 32) {
    const int i = base + lane;
    const bool ok = i < m && A.lp[f + 1 + i] < A.rp[f + i];
    const unsigned b = __ballot_sync(0xffffffffu, ok);
    if (b == 0xffffffffu) {
      t += 32;
      continue;
    }
    t += __ffs(~b) - 1;
    break;
  }
  for (int i = lane; i < t; i += 32) ord_swap(A, A.lp[f + 1 + i], A.rp[f + i]);
  int cut;
  if (t == 0) {
    cut = nL > 0 ? A.lp[f + 1] : l;
  } else {
    cut = A.rp[f + t - 1];
    if (t < nL) {
      const int nextL = A.lp[f + 1 + t];
      if (nextL < cut) cut = nextL;
    }
  }
  __syncwarp();
  return cut;
}

// The same for a range of at most 32 elements (60 % of all partitions of a tree): one element per lane, in registers;
// median, pivot and the exchange of the swapped pairs by shuffles; only the two stop lists go through shared memory.
// A third of the instructions of the general form - the partitions are bound by the length of their dependent
// instruction chain (narrow levels of the tree) or by instruction issue (wide levels).
__device__ __forceinline__ int ord_warp_partition_small(const OrdArrays& A, int f, int l, int lane) {
  const int len = l - f;   // 17 .. 32
  const bool in = lane < len;
  int sc = in ? A.sc[f + lane] : 0;
  uint32_t ra = in ? A.ra[f + lane] : 0u;
  const int mid = len >> 1;
  int pick;
  {  // __move_median_to_first(first, first + 1, mid, last - 1)
    const int sa = __shfl_sync(0xffffffffu, sc, 1), sb = __shfl_sync(0xffffffffu, sc, mid),
              s_c = __shfl_sync(0xffffffffu, sc, len - 1);
    if (sa > sb) pick = (sb > s_c) ? mid : ((sa > s_c) ? len - 1 : 1);
    else pick = (sa > s_c) ? 1 : ((sb > s_c) ? len - 1 : mid);
  }
  const int s0 = __shfl_sync(0xffffffffu, sc, 0), pv = __shfl_sync(0xffffffffu, sc, pick);
  const uint32_t r0 = __shfl_sync(0xffffffffu, ra, 0), rp = __shfl_sync(0xffffffffu, ra, pick);
  if (lane == 0) {
    sc = pv;
    ra = rp;
  } else if (lane == pick) {
    sc = s0;
    ra = r0;
  }
  const bool isL = in && lane >= 1 && !(sc > pv);   // "while (comp(*first, pivot)) ++first" stops here
  const bool isR = in && !(pv > sc);                // "while (comp(pivot, *last)) --last" stops here (lane 0 = sentinel)
  const unsigned bL = __ballot_sync(0xffffffffu, isL), bR = __ballot_sync(0xffffffffu, isR);
  const int nL = __popc(bL), nR = __popc(bR);
  const unsigned lt = (1u << lane) - 1u;
  const int rL = __popc(bL & lt);                       // rank among the left stops, ascending
  const int rR = __popc(bR & ~(lt | (1u << lane)));     // rank among the right stops, descending
  if (isL) A.lp[f + 1 + rL] = lane;
  if (isR) A.rp[f + rR] = lane;
  __syncwarp();
  // pair j = (j-th left stop, j-th right stop) is swapped while L_j < R_j; an element equal to the pivot is a stop of
  // both kinds but can belong to one swapped pair only
  int partner = lane;
  bool leftOfPair = false;
  if (isL && rL < nR) {
    const int p = A.rp[f + rL];
    if (lane < p) {
      partner = p;
      leftOfPair = true;
    }
  }
  if (!leftOfPair && isR && rR < nL) {
    const int p = A.lp[f + 1 + rR];
    if (p < lane) partner = p;
  }
  const int t = __popc(__ballot_sync(0xffffffffu, leftOfPair));
  const int nsc = __shfl_sync(0xffffffffu, sc, partner);
  const uint32_t nra = __shfl_sync(0xffffffffu, ra, partner);
  if (in) {
    A.sc[f + lane] = nsc;
    A.ra[f + lane] = nra;
  }
  int cut;
  if (t == 0) {
    cut = nL > 0 ? __ffs(bL) - 1 : len;
  } else {
    cut = A.rp[f + t - 1];
    if (t < nL) {
      const int nextL = A.lp[f + 1 + t];
      if (nextL < cut) cut = nextL;
    }
  }
  __syncwarp();
  return f + cut;
}

// The same partition by the whole CTA, for the long ranges at the top of the recursion (one warp walking 7 000
// elements twice was most of the kernel): the warps classify contiguous chunks, a scan over the warps places the
// stop lists, and because the left stops ascend while the right stops descend, "L_i < R_i" holds for a prefix of the
// pairs, so the number of swaps is a plain count. Called by every thread; returns the cut to every thread.
struct OrdBlockScratch { int cntL[32], cntR[32], offL[32], offR[32], nL, nR, t; };

__device__ __forceinline__ int ord_block_partition(const OrdArrays& A, int f, int l, OrdBlockScratch& S) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {  // __move_median_to_first(first, first + 1, mid, last - 1)
    const int a = f + 1, b = f + (l - f) / 2, c = l - 1;
    const int sa = A.sc[a], sb = A.sc[b], s_c = A.sc[c];
    int pick;
    if (sa > sb) pick = (sb > s_c) ? b : ((sa > s_c) ? c : a);
    else pick = (sa > s_c) ? a : ((sb > s_c) ? c : b);
    ord_swap(A, f, pick);
    S.t = 0;
  }
  __syncthreads();
  const int pv = A.sc[f];
  const int len = l - f;
  const unsigned lt = (1u << lane) - 1u;
  const int chunk = (((len + 31) / 32 + 31) / 32) * 32;   // indices per warp, a multiple of 32
  const int i0 = warp * chunk, i1 = min(len, i0 + chunk);
  int nL = 0, nR = 0;
  for (int base = i0; base < i1; base += 32) {
    const int i = base + lane;
    const int pL = f + 1 + i, pR = l - 1 - i;
    const bool isL = i < i1 && pL < l && !(A.sc[pL] > pv);
    const bool isR = i < i1 && !(pv > A.sc[pR]);
    nL += __popc(__ballot_sync(0xffffffffu, isL));
    nR += __popc(__ballot_sync(0xffffffffu, isR));
  }
  if (lane == 0) {
    S.cntL[warp] = nL;
    S.cntR[warp] = nR;
  }
  __syncthreads();
  if (warp == 0) {
    const int cl = S.cntL[lane], cr = S.cntR[lane];
    int il = cl, ir = cr;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int tl = __shfl_up_sync(0xffffffffu, il, d), tr = __shfl_up_sync(0xffffffffu, ir, d);
      if (lane >= d) {
        il += tl;
        ir += tr;
      }
    }
    S.offL[lane] = il - cl;
    S.offR[lane] = ir - cr;
    if (lane == 31) {
      S.nL = il;
      S.nR = ir;
    }
  }
  __syncthreads();
  int oL = S.offL[warp], oR = S.offR[warp];
  for (int base = i0; base < i1; base += 32) {
    const int i = base + lane;
    const int pL = f + 1 + i, pR = l - 1 - i;
    const bool isL = i < i1 && pL < l && !(A.sc[pL] > pv);
    const bool isR = i < i1 && !(pv > A.sc[pR]);
    const unsigned bL = __ballot_sync(0xffffffffu, isL), bR = __ballot_sync(0xffffffffu, isR);
    if (isL) A.lp[f + 1 + oL + __popc(bL & lt)] = pL;
    if (isR) A.rp[f + oR + __popc(bR & lt)] = pR;
    oL += __popc(bL);
    oR += __popc(bR);
  }
  __syncthreads();
  const int NL = S.nL, NR = S.nR;
  const int m = NL < NR ? NL : NR;
  int mine = 0;
  for (int i = tid; i < m; i += kOrdThreads) mine += A.lp[f + 1 + i] < A.rp[f + i] ? 1 : 0;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
  if (lane == 0 && mine) atomicAdd(&S.t, mine);
  __syncthreads();
  const int t = S.t;
  int cut;
  if (t == 0) {
    cut = NL > 0 ? A.lp[f + 1] : l;
  } else {
    cut = A.rp[f + t - 1];
    if (t < NL) {
      const int nextL = A.lp[f + 1 + t];
      if (nextL < cut) cut = nextL;
    }
  }
  // (no barrier: the swaps move elements of A.sc / A.ra, the stop lists read above hold positions)
  for (int i = tid; i < t; i += kOrdThreads) ord_swap(A, A.lp[f + 1 + i], A.rp[f + i]);
  __syncthreads();
  return cut;
}

constexpr int kOrdCoopLen = 1024;   // ranges at least this long are partitioned by the whole CTA
constexpr int kShareSmMinSlots = 8;  // launches of at least this many images use the 48-register builds of the one-CTA-per-image kernels

// kRegs = 48 for batches: the kernel then leaves a quarter of the register file of its SM (it is alone there: 1 024
// threads, 200 KB of shared memory) to CTAs of the GPU-wide kernels of the other batches in flight, which use the issue
// slots this latency-bound kernel leaves idle (+4.5 % batch throughput); 64 for a multiframe alone (4 % less latency).
template <int kRegs>
__global__ void __maxnreg__(kRegs)
order_candidates_kernel(const Slot* __restrict__ slots, int W, int H, bool denseScores) {
  const Slot& slot = slots[blockIdx.x];
  const int32_t* __restrict__ scores = slot.scores;
  const uint32_t* __restrict__ bitmap = slot.bitmap;
  uint64_t* __restrict__ cand = slot.cand;
  int* __restrict__ counters = slot.counters;
  uint8_t* __restrict__ gscratch = slot.order_scratch;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_warp[32];
  __shared__ int s_n, s_qcount[2];
  __shared__ unsigned s_cut[kCandCap / 32 + 1];
  __shared__ OrdBlockScratch s_blk;
  __shared__ int2 s_big[2][kCandCap / kOrdCoopLen + 1];   // ranges of the current / next level long enough for the CTA
  __shared__ int s_bigcount[2];   // bit i: a partition cut (or the start) lies at position i
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef OKC_ORDER_TRACE
  long long tr[10] = {}, t0 = clock64();
#define ORD_T(i) do { __syncthreads(); if (tid == 0) tr[i] += clock64() - t0; t0 = clock64(); } while (0)
#else
#define ORD_T(i) do { } while (0)
#endif

  // ---- A. candidate count and raster-ordered offsets from the bitmap ----
  const int wordsPerRow = harris_strips(W) * kStripWords;
  const int words = H * wordsPerRow;
  // every warp owns a contiguous block of bitmap words and reads it 32 words at a time (coalesced)
  const int chunk = (((words + 31) / 32 + 31) / 32) * 32;   // words per warp, a multiple of 32
  const int w0 = warp * chunk, w1 = min(words, w0 + chunk);
  // Usual image sizes (<= 16 384 words = 491 520 pixels): a thread keeps 16 consecutive words (64 bytes, 4 independent
  // 16-byte loads) in registers for both passes, and its candidates are consecutive in raster order - one scan per
  // warp instead of one per 32 words (the per-word-row scans and divisions made this phase instruction-issue bound:
  // 15 000 cycles for 13 440 words). Larger images: the strided two-pass form.
  constexpr int kLaneWords = 16;
  const bool fast = words <= kOrdThreads * kLaneWords && (reinterpret_cast<uintptr_t>(bitmap) & 15) == 0;
  unsigned wreg[kLaneWords];
  int mine = 0;
  if (fast) {
#pragma unroll
    for (int u = 0; u < kLaneWords / 4; ++u) {
      const int w = tid * kLaneWords + 4 * u;   // (words is a multiple of kStripWords = 4)
      const uint4 v = w < words ? __ldg(reinterpret_cast<const uint4*>(bitmap + w)) : make_uint4(0, 0, 0, 0);
      wreg[4 * u] = v.x;
      wreg[4 * u + 1] = v.y;
      wreg[4 * u + 2] = v.z;
      wreg[4 * u + 3] = v.w;
    }
#pragma unroll
    for (int j = 0; j < kLaneWords; ++j) mine += __popc(wreg[j]);
  } else {
    for (int w = w0 + lane; w < w1; w += 32) mine += __popc(bitmap[w]);
  }
  int incl = mine;
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = s_warp[lane], iv = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, iv, d);
      if (lane >= d) iv += u;
    }
    s_warp[lane] = iv - v;
    if (lane == 31) s_n = iv;
  }
  __syncthreads();
  int total = s_n;
  ORD_T(6);
  if (total > kCandCap) {
    if (tid == 0) {
      counters[0] = kCandCap;
      counters[3] = 1;  // overflow: reported by the C ABI as OKC_ERR_CAPACITY
    }
    return;
  }
  const int n = total;
  if (tid == 0) counters[0] = n;
  if (n == 0) return;

  // The body is instantiated twice - arrays in shared memory (n <= kOrdSmemCap) or in the slot's global scratch - so
  // that the shared-memory instance is compiled to LDS / STS with 32-bit addresses: through one generic pointer for
  // both, every access of the partitions' dependent load -> compare -> store chains was a generic LD.E / ST.E.
  auto body = [&](auto inSmemTag) {
  OrdArrays A;
  constexpr bool inSmem = decltype(inSmemTag)::value;
  {
    uint8_t* base = inSmem ? smem : gscratch;
    const size_t cap = inSmem ? kOrdSmemCap : kCandCap;
    const size_t qcap = inSmem ? kOrdQueueSmem : kOrdQueueGlobal;
    A.ra = reinterpret_cast<uint32_t*>(base);
    A.q[0] = reinterpret_cast<int2*>(base + cap * 4);
    A.q[1] = A.q[0] + qcap;
    uint8_t* u = base + cap * 4 + 2 * qcap * 8;
    A.sc = reinterpret_cast<int*>(u);
    A.lp = A.sc + cap;
    A.rp = A.lp + cap;
  }
  for (int w = tid; w <= (n >> 5); w += kOrdThreads) s_cut[w] = w == 0 ? 1u : 0u;
  if (fast) {
    // (thread order = raster order: `incl` is the inclusive scan of `mine` over the warp's lanes, from the first pass)
    int pos = s_warp[warp] + incl - mine;
    const int wFirst = tid * kLaneWords;
    int y = wFirst / wordsPerRow, rem = wFirst - y * wordsPerRow;
    int segAt = 0;
#pragma unroll
    for (int j = 0; j < kLaneWords; ++j) {
      unsigned bits = wreg[j];
      // compact scores (harris_nms_kernel<.., false>): a (row, strip) segment = 4 words; a candidate's score is at its
      // rank inside the segment. Stored first: WHERE the score is; the scores follow in one pass of independent loads
      if ((j & (kStripWords - 1)) == 0) segAt = ((wFirst + j) / kStripWords) * kStripCols;
      const int xb = (rem / kStripWords) * kStripCols - 4 + ((rem % kStripWords) << 5);
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t raster = (uint32_t)(y * W + xb + b);
        A.sc[pos] = denseScores ? (int)raster : segAt++;
        A.ra[pos] = raster;
        ++pos;
      }
      if (++rem == wordsPerRow) {
        rem = 0;
        ++y;
      }
    }
  } else {
    int base = s_warp[warp];   // candidates in front of this warp's block
    for (int wb = w0; wb < w1; wb += 32) {
      const int w = wb + lane;
      unsigned bits = w < w1 ? bitmap[w] : 0u;
      const int c = __popc(bits);
      int inc = c;
      for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += v;
      }
      int pos = base + inc - c;
      base += __shfl_sync(0xffffffffu, inc, 31);
      const int y = w / wordsPerRow, rem = w - y * wordsPerRow;
      const int xb = (rem / kStripWords) * kStripCols - 4 + ((rem % kStripWords) << 5);
      // compact scores (harris_nms_kernel<.., false>): the 4 words of a (row, strip) segment sit in 4 adjacent lanes
      // (wb is a multiple of 32), a candidate's score is at its rank inside the segment
      const int excl = inc - c;
      int segRank = excl - __shfl_sync(0xffffffffu, excl, lane & ~(kStripWords - 1));
      // first WHERE the score is (its index in the score buffer), then all scores in one pass of independent loads:
      // a load per candidate inside this loop was a chain of L2 round trips (2/3 of the time of this phase)
      int segAt = (w / kStripWords) * kStripCols + segRank;
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t raster = (uint32_t)(y * W + xb + b);
        A.sc[pos] = denseScores ? (int)raster : segAt++;
        A.ra[pos] = raster;
        ++pos;
      }
    }
  }
  __syncthreads();
  ORD_T(7);
  for (int i0 = tid; i0 < n; i0 += 8 * kOrdThreads) {
    int at[8], v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) at[u] = i0 + u * kOrdThreads < n ? A.sc[i0 + u * kOrdThreads] : 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldg(scores + at[u]);
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i0 + u * kOrdThreads < n) A.sc[i0 + u * kOrdThreads] = v[u];
  }
  if (tid == 0) {
    s_qcount[0] = s_qcount[1] = s_bigcount[0] = s_bigcount[1] = 0;
    if (n >= kOrdCoopLen) s_big[0][s_bigcount[0]++] = make_int2(0, n);
    else if (n > 16) A.q[0][s_qcount[0]++] = make_int2(0, n);
  }
  __syncthreads();
  ORD_T(0);

  // ---- B. introsort partition phase, level by level ----
  int depthLimit = 2 * (31 - __clz(n));
  int cur = 0;
  while (true) {
    const int cnt = s_qcount[cur], cntBig = s_bigcount[cur];
    if (cnt == 0 && cntBig == 0) break;
    // a range that is split further goes to the next level's queue: the CTA's when it is long, else the warps'
    auto push = [&](int a, int b) {
      if (b - a >= kOrdCoopLen) s_big[cur ^ 1][atomicAdd(&s_bigcount[cur ^ 1], 1)] = make_int2(a, b);
      else if (b - a > 16) A.q[cur ^ 1][atomicAdd(&s_qcount[cur ^ 1], 1)] = make_int2(a, b);
    };
#ifdef OKC_ORDER_TRACE
    long long tl0 = clock64(), tl1 = 0;
    int trMax = 0;
    for (int s = 0; s < cnt; ++s) trMax = max(trMax, A.q[cur][s].y - A.q[cur][s].x);
#endif
    for (int s = 0; s < cntBig; ++s) {   // the long ranges, by the whole CTA (uniform loop: every thread takes part)
      const int2 r = s_big[cur][s];
      if (depthLimit == 0) {
        if (tid == 0) ord_heap_sort(A, r.x, r.y);
        continue;
      }
      const int cut = ord_block_partition(A, r.x, r.y, s_blk);
      if (tid == 0) {
        atomicOr(&s_cut[cut >> 5], 1u << (cut & 31));
        push(r.x, cut);
        push(cut, r.y);
      }
    }
#ifdef OKC_ORDER_TRACE
    __syncthreads();
    tl1 = clock64();
#endif
    for (int s = warp; s < cnt; s += kOrdThreads / 32) {
      const int2 r = A.q[cur][s];
      if (depthLimit == 0) {
        if (lane == 0) ord_heap_sort(A, r.x, r.y);
        continue;
      }
      const int cut = r.y - r.x <= 32 ? ord_warp_partition_small(A, r.x, r.y, lane) : ord_warp_partition(A, r.x, r.y, lane);
      if (lane == 0) {
        atomicOr(&s_cut[cut >> 5], 1u << (cut & 31));
        push(r.x, cut);
        push(cut, r.y);
      }
    }
    __syncthreads();
#ifdef OKC_ORDER_TRACE
    if (tid == 0 && n == 7517) printf("  level: %d long (%lld cycles), %d warp ranges, longest %d (%lld cycles)\n", cntBig, tl1 - tl0, cnt, trMax, clock64() - tl1);
#endif
    if (tid == 0) s_qcount[cur] = s_bigcount[cur] = 0;
    cur ^= 1;
    --depthLimit;
    __syncthreads();
  }
  ORD_T(1);
  // ---- C. __final_insertion_sort = a stable insertion sort of the whole array. The partitions leave it as a chain
  //         of segments between consecutive cuts - at most 16 elements each, or heap-sorted - where no element of a
  //         later segment belongs in front of an earlier one, so the stable sort of the array is the stable sort of
  //         every segment ----
  // Every element finds its segment [start, end) in the cut bitmap and its rank inside it - the elements with a larger
  // score plus the equal ones in front of it - i.e. its place after a stable sort. Segments longer than 16 elements
  // only come from the heapsort path and are sorted already.
  for (int i = tid; i < n; i += kOrdThreads) {
    int start, end = n;
    {
      int w = i >> 5;
      unsigned bits = s_cut[w] & (0xffffffffu >> (31 - (i & 31)));   // cuts at or before i in its word
      while (!bits) bits = s_cut[--w];                                // (bit 0 of word 0 is always set)
      start = (w << 5) + 31 - __clz(bits);
      w = i >> 5;
      bits = (i & 31) == 31 ? 0u : (s_cut[w] & (0xffffffffu << ((i & 31) + 1)));   // cuts behind i
      for (;;) {
        if (bits) {
          end = min(n, (w << 5) + __ffs(bits) - 1);
          break;
        }
        if (++w > (n >> 5)) break;
        bits = s_cut[w];
      }
    }
    const int vs = A.sc[i];
    int pos = i;
    if (end - start <= 16) {
      int ahead = 0;
      for (int j = start; j < end; ++j) {
        const int sj = A.sc[j];
        ahead += (sj > vs || (sj == vs && j < i)) ? 1 : 0;
      }
      pos = start + ahead;
    }
    A.lp[pos] = vs;
    A.rp[pos] = (int)A.ra[i];
  }
  __syncthreads();
  ORD_T(3);
  for (int i = tid; i < n; i += kOrdThreads) cand[i] = cand_key(A.lp[i], (uint32_t)A.rp[i]);
#ifdef OKC_ORDER_TRACE
  ORD_T(4);
  if (tid == 0) printf("  gather: count %lld place %lld scores+rest %lld\n", tr[6], tr[7], tr[0]);
  if (tid == 0) printf("order trace (cycles): n %d gather %lld partition %lld segments %lld write %lld\n", n, tr[0], tr[1], tr[3], tr[4]);
#endif
  };
  if (n <= kOrdSmemCap) body(std::true_type{});
  else body(std::false_type{});
}

constexpr size_t kOrdSmemBytes = (size_t)kOrdSmemCap * 4 + 2 * (size_t)kOrdQueueSmem * 8 + (size_t)kOrdSmemCap * 12;
size_t order_scratch_bytes() {
  return (size_t)kCandCap * 4 + 2 * (size_t)kOrdQueueGlobal * 8 + (size_t)kCandCap * 12 + (size_t)kCandCap * 8 + 256;
}

void launch_order_candidates(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s) {
  // the attribute is per device (a process may hold contexts on several)
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    cudaFuncSetAttribute(order_candidates_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kOrdSmemBytes);
    cudaFuncSetAttribute(order_candidates_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kOrdSmemBytes);
    configured[dev & 63] = true;
  }
  if (n >= kShareSmMinSlots) order_candidates_kernel<48><<<n, kOrdThreads, kOrdSmemBytes, s>>>(slots, a.width, a.height, a.dense_scores);
  else order_candidates_kernel<64><<<n, kOrdThreads, kOrdSmemBytes, s>>>(slots, a.width, a.height, a.dense_scores);
}

// ----------------------------------------------------------------------------------------------
// greedy uniformity + sub-pixel refinement: one CTA per image
// ----------------------------------------------------------------------------------------------
constexpr int kSelThreads = 256;
constexpr int kSelCells = (31 * 31 + kSelThreads - 1) / kSelThreads;   // stamp cells per thread

// Subpixel2D, scale-space-layer-inl.h:568-701. Returns only the deltas (the refined peak value is
// computed and discarded by the caller in the reference, :403-411).
__device__ void subpixel2d(double s_0_0, double s_0_1, double s_0_2, double s_1_0, double s_1_1, double s_1_2,
                           double s_2_0, double s_2_1, double s_2_2, float& delta_x, float& delta_y) {
  double tmp1 = s_0_0 + s_0_2 - 2 * s_1_1 + s_2_0 + s_2_2;
  double coeff1 = 3 * (tmp1 + s_0_1 - ((s_1_0 + s_1_2) * 0.5) + s_2_1);   // (x / 2.0: exact either way)
  double coeff2 = 3 * (tmp1 - ((s_0_1 + s_2_1) * 0.5) + s_1_0 + s_1_2);
  double tmp2 = s_0_2 - s_2_0;
  double tmp3 = (s_0_0 + tmp2 - s_2_2);
  double tmp4 = tmp3 - 2 * tmp2;
  double coeff3 = -3 * (tmp3 + s_0_1 - s_2_1);
  double coeff4 = -3 * (tmp4 + s_1_0 - s_1_2);
  double coeff5 = (s_0_0 - s_0_2 - s_2_0 + s_2_2) * 0.25;                  // (x / 4.0: exact either way)
  double H_det = 4 * coeff1 * coeff2 - coeff5 * coeff5;
  if (H_det == 0) {
    delta_x = 0.0f;
    delta_y = 0.0f;
    return;
  }
  if (!(H_det > 0 && coeff1 < 0)) {
    // maximum at one of the 4 patch corners; note the int truncation of the sums (:598-620)
    int tmp_max = (int)(coeff3 + coeff4 + coeff5);
    delta_x = 1.0f;
    delta_y = 1.0f;
    int tmp = (int)(-coeff3 + coeff4 - coeff5);
    if (tmp > tmp_max) {
      tmp_max = tmp;
      delta_x = -1.0f;
      delta_y = 1.0f;
    }
    tmp = (int)(coeff3 - coeff4 - coeff5);
    if (tmp > tmp_max) {
      tmp_max = tmp;
      delta_x = 1.0f;
      delta_y = -1.0f;
    }
    tmp = (int)(-coeff3 - coeff4 + coeff5);
    if (tmp > tmp_max) {
      tmp_max = tmp;
      delta_x = -1.0f;
      delta_y = -1.0f;
    }
    return;
  }
  delta_x = (float)(2 * coeff2 * coeff3 - coeff4 * coeff5) / (float)(-H_det);
  delta_y = (float)(2 * coeff1 * coeff4 - coeff3 * coeff5) / (float)(-H_det);
  bool tx = false, tx_ = false, ty = false, ty_ = false;
  if (delta_x > 1.0f) tx = true;
  else if (delta_x < -1.0f) tx_ = true;
  if (delta_y > 1.0f) ty = true;
  if (delta_y < -1.0f) ty_ = true;
  if (tx || tx_ || ty || ty_) {
    double coeff6 = -(s_0_0 + s_0_2 - ((s_1_0 + s_0_1 + s_1_2 + s_2_1) / 2.0) - 5 * s_1_1 + s_2_0 + s_2_2) / 2.01;
    float delta_x1 = 0.0f, delta_x2 = 0.0f, delta_y1 = 0.0f, delta_y2 = 0.0f;
    if (tx) {
      delta_x1 = 1.0f;
      delta_y1 = -(float)(coeff4 + coeff5) / (float)(2 * coeff2);
      if (delta_y1 > 1.0f) delta_y1 = 1.0f;
      else if (delta_y1 < -1.0f) delta_y1 = -1.0f;
    } else if (tx_) {
      delta_x1 = -1.0f;
      delta_y1 = -(float)(coeff4 - coeff5) / (float)(2 * coeff2);
      if (delta_y1 > 1.0f) delta_y1 = 1.0f;
      else if (delta_y1 < -1.0f) delta_y1 = -1.0f;
    }
    if (ty) {
      delta_y2 = 1.0f;
      delta_x2 = -(float)(coeff3 + coeff5) / (float)(2 * coeff1);
      if (delta_x2 > 1.0f) delta_x2 = 1.0f;
      else if (delta_x2 < -1.0f) delta_x2 = -1.0f;
    } else if (ty_) {
      delta_y2 = -1.0f;
      delta_x2 = -(float)(coeff3 - coeff5) / (float)(2 * coeff1);
      if (delta_x2 > 1.0f) delta_x2 = 1.0f;
      else if (delta_x2 < -1.0f) delta_x2 = -1.0f;
    }
    // double expression, narrowed to float on assignment (:676-681)
    float max1 = (float)((coeff1 * delta_x1 * delta_x1 + coeff2 * delta_y1 * delta_y1 + coeff3 * delta_x1 +
                          coeff4 * delta_y1 + coeff5 * delta_x1 * delta_y1 + coeff6) / 18.0);
    float max2 = (float)((coeff1 * delta_x2 * delta_x2 + coeff2 * delta_y2 * delta_y2 + coeff3 * delta_x2 +
                          coeff4 * delta_y2 + coeff5 * delta_x2 * delta_y2 + coeff6) / 18.0);
    if (max1 > max2) {
      delta_x = delta_x1;
      delta_y = delta_x1;  // sic (:684-686)
    } else {
      delta_x = delta_x2;
      delta_y = delta_x2;  // sic (:688-690)
    }
  }
}

// The 3x3 block of Harris scores around pixel (u, v), recomputed from the image with the arithmetic of
// harris_nms_kernel (same integers, so the same values): 7x7 pixels -> 5x5 gradient products -> 3x5 horizontally
// smoothed -> 3x3 scores; 0 outside rows 2..H-3 / columns 2..W-3 as in the map. out[3 * (dv + 1) + (du + 1)].
__device__ __forceinline__ void harris_ring_scores(const uint8_t* __restrict__ img, int W, int H, int u, int v,
                                                   int out[9]) {
  int h[7][5], sm[7][5];   // per pixel row v-3..v+3, columns u-2..u+2: I[x-1] - I[x+1] and 8 * (3, 10, 3)-sums
#pragma unroll
  for (int pr = 0; pr < 7; ++pr) {
    const int r = v - 3 + pr;
    const int rr = r < 0 ? 0 : (r >= H ? H - 1 : r);
    int px[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int xx = u - 3 + i;
      px[i] = (xx >= 0 && xx < W) ? (int)__ldg(img + (size_t)rr * W + xx) : 0;
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      h[pr][i] = px[i] - px[i + 2];
      sm[pr][i] = 24 * (px[i] + px[i + 2]) + 80 * px[i + 1];
    }
  }
  int sxx[5][3], sxy[5][3], syy[5][3];   // horizontally smoothed products: gradient rows v-2..v+2, columns u-1..u+1
#pragma unroll
  for (int g = 0; g < 5; ++g) {          // gradient row v-2+g reads pixel rows g, g+1, g+2 of the 7
    int pxx[5], pxy[5], pyy[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const int dx = 24 * (h[g][i] + h[g + 2][i]) + 80 * h[g + 1][i];
      const int dy = sm[g][i] - sm[g + 2][i];
      pxx[i] = (dx * dx) >> 16;
      pxy[i] = (dx * dy) >> 16;
      pyy[i] = (dy * dy) >> 16;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      sxx[g][i] = pxx[i] + 2 * pxx[i + 1] + pxx[i + 2];
      sxy[g][i] = pxy[i] + 2 * pxy[i + 1] + pxy[i + 2];
      syy[g][i] = pyy[i] + 2 * pyy[i + 1] + pyy[i + 2];
    }
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int y = v - 1 + j;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int x = u - 1 + i;
      const int a = (sxx[j][i] + 2 * sxx[j + 1][i] + sxx[j + 2][i]) >> 4;
      const int b = (syy[j][i] + 2 * syy[j + 1][i] + syy[j + 2][i]) >> 4;
      const int d = (sxy[j][i] + 2 * sxy[j + 1][i] + sxy[j + 2][i]) >> 4;
      const int t = (a + b) >> 1;
      const int sc = a * b - d * d - (int)((unsigned)(t * t) / 5u);
      out[3 * j + i] = (y >= 2 && y <= H - 3 && x >= 2 && x <= W - 3) ? sc : 0;
    }
  }
}

template <bool kOccInSmem>
__global__ void __launch_bounds__(kSelThreads, 1)
select_keypoints_kernel(const Slot* __restrict__ slots, const float* __restrict__ lut_global, int W, int H,
                        float scaling, unsigned long long maxKpt, int maxOut, int smemBytes, double kpScale,
                        double kpOffset, int octave) {
  const Slot& slot = slots[blockIdx.x];
  uint64_t* __restrict__ cand = slot.cand;
  int* __restrict__ counters = slot.counters;
  uint8_t* __restrict__ occ_global = slot.occupancy;
  okc_keypoint* __restrict__ out = slot.kp_detect;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ float lut[31 * 31];
  __shared__ int s_off[kSelThreads];
  __shared__ float s_nsc[kSelThreads];
  __shared__ uint64_t s_key[kSelThreads];   // (a global re-read of the key by one thread would stall every barrier)
  __shared__ int s_warpcnt[32], s_warpoff[32], s_total;

  const int tid = threadIdx.x;
  int n = counters[0];
  if (n > kCandCap) n = kCandCap;
  if (counters[3] == 1) n = 0;   // candidate overflow: cand[] was not written (reported as OKC_ERR_CAPACITY)
  for (int i = tid; i < 31 * 31; i += kSelThreads) lut[i] = lut_global[i];
  if (n == 0) {
    if (tid == 0) counters[1] = 0;
    return;
  }

  // cand[0..n) arrives in the reference's std::sort order (order_candidates_kernel)

  // ---- occupancy image: only [0, occRows) x [0, occCols) of the reference's (H*ceil+32)x(W*ceil+32) image is
  //      ever addressed (uniformity-enforcement-inl.h:75-86); keep it in shared memory when it fits ----
  const int occRows = (int)((float)(H - 3) * scaling + 16.0f) + 16 + 1;
  const int occStride = (((int)((float)(W - 3) * scaling + 16.0f) + 16 + 1) + 15) & ~15;
  const size_t occBytes = (size_t)occRows * occStride;
  uint8_t* occ = kOccInSmem ? smem : occ_global;   // (shared-memory instructions instead of generic ones)
  for (size_t i = (size_t)tid * 16; i < occBytes; i += (size_t)kSelThreads * 16)
    *reinterpret_cast<uint4*>(occ + i) = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const float maxScore = (float)cand_score(cand[0]);
  float myLut[kSelCells];
  int stampOff[kSelCells];
#pragma unroll
  for (int c = 0; c < kSelCells; ++c) {
    const int cell = tid + c * kSelThreads;
    const int ly = cell / 31, lx = cell - ly * 31;
    myLut[c] = cell < 31 * 31 ? lut[cell] : 0.0f;
    stampOff[c] = (ly - 15) * occStride + (lx - 15);
  }
  const int lane = tid & 31, warp = tid >> 5;
  int kept = 0;
  bool done = false;

  // The loop is sequential in the reference. Occupancy only grows (saturating add of non-negative
  // stamps), so a candidate that fails "nsc1 < occupancy" once fails forever: each chunk of 1024
  // candidates is tested in parallel against the current occupancy, the survivors are compacted in
  // order, and only those are visited one by one (re-tested at visit time, exactly like the reference).
  for (int base = 0; base < n && !done; base += kSelThreads) {
    const int i = base + tid;
    bool alive = false;
    int off = 0;
    float nsc1 = 0.0f;
    uint64_t key = 0;
    if (i < n) {
      key = cand[i];
      const int score = cand_score(key);
      const uint32_t raster = (uint32_t)key;
      const int y = raster / W, x = raster - y * W;
      const int cy = (int)((float)y * scaling + 16.0f);
      const int cx = (int)((float)x * scaling + 16.0f);
      off = cy * occStride + cx;
      nsc1 = sqrtf(sqrtf((float)score / maxScore)) * 255.0f;
      alive = !(nsc1 < (float)occ[off]);
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, alive);
    if (lane == 0) s_warpcnt[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
      int c = lane < kSelThreads / 32 ? s_warpcnt[lane] : 0, incl = c;
      for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      s_warpoff[lane] = incl - c;
      if (lane == 31) s_total = incl;
    }
    __syncthreads();
    if (alive) {
      const int pos = s_warpoff[warp] + __popc(ballot & ((1u << lane) - 1u));
      s_off[pos] = off;
      s_nsc[pos] = nsc1;
      s_key[pos] = key;
    }
    __syncthreads();
    const int total = s_total;
    for (int j = 0; j < total; ++j) {
      const int o = s_off[j];
      const float v1 = s_nsc[j];
      if (v1 < (float)occ[o]) continue;
      const float nsc = 0.99f * v1;
      __syncthreads();  // everyone has read the occupancy before the stamp changes it
      if (nsc == nsc) {  // a NaN stamp adds 0 everywhere (cvttss2si + byte truncation in the reference)
        // the thread's cells are distinct: all loads first, then all stores (byte stores would otherwise order the loads)
        int old[kSelCells];
#pragma unroll
        for (int c = 0; c < kSelCells; ++c)
          if (tid + c * kSelThreads < 31 * 31) old[c] = (int)occ[o + stampOff[c]];
#pragma unroll
        for (int c = 0; c < kSelCells; ++c) {
          if (tid + c * kSelThreads < 31 * 31) {
            const int sum = old[c] + ((int)ceilf(myLut[c] * nsc) & 0xff);
            occ[o + stampOff[c]] = (uint8_t)(sum > 255 ? 255 : sum);
          }
        }
      }
      if (tid == 0 && kept < maxOut) cand[kCandCap + kept] = s_key[j];  // kept list lives behind the keys
      ++kept;
      __syncthreads();
      if ((unsigned long long)kept == maxKpt) {
        done = true;
        break;
      }
    }
    __syncthreads();
  }

  // ---- sub-pixel refinement -> cv::KeyPoint (scale-space-layer-inl.h:394-420, layer 0: _scale 1, _offset 0) ----
  int nOut = kept < maxOut ? kept : maxOut;
  if (tid == 0) {
    counters[1] = nOut;
    if (kept > maxOut) counters[3] = 2;
  }
  for (int i = tid; i < nOut; i += kSelThreads) {
    const uint64_t key = cand[kCandCap + i];
    const int score = cand_score(key);
    const uint32_t raster = (uint32_t)key;
    const int v = raster / W, u = raster - v * W;
    int c[9];
    harris_ring_scores(slot.image, W, H, u, v, c);
    float dx, dy;
    subpixel2d((double)c[0], (double)c[1], (double)c[2], (double)c[3], (double)c[4], (double)c[5], (double)c[6],
               (double)c[7], (double)c[8], dx, dy);
    okc_keypoint kp;
    // _scale * ((it->x + delta_x) + _offset): (float + float) widened, the rest in double, narrowed on assignment
    kp.x = (float)(kpScale * ((double)((float)u + dx) + kpOffset));
    kp.y = (float)(kpScale * ((double)((float)v + dy) + kpOffset));
    kp.size = (float)(kpScale * 12.0);
    kp.angle = -1.0f;
    kp.response = (float)score;
    kp.octave = octave;
    kp.class_id = -1;
    out[i] = kp;
  }
}

// ----------------------------------------------------------------------------------------------
// Dependency form of the greedy loop (default): no occupancy image, no 31x31 stamps, no barrier per decision.
//
// The reference visits the sorted candidates one by one; a candidate's decision reads the occupancy at ONE cell, and
// that byte is min(255, sum of the stamp values of the ACCEPTED earlier candidates whose own cell lies within +-15
// cells) - saturating adds of non-negative values commute. So a candidate only depends on the lower-position
// candidates within +-15 cells ("blockers"), and once those are decided its own decision is final.
//   window  : up to 1024 candidates in sorted order, one per thread. A new candidate is first tested against a lower
//             bound of the occupancy (one read), then gathers its exact sum from the points accepted in earlier
//             windows (CSR by spatial bin) and is dropped if that already rejects it (the sum only grows, so the
//             reference would reject it at its turn).
//   blockers: the window entries are binned (CSR over bins of >= 16 cells); every entry lists its blockers from the
//             3x3 bins around it (an entry with more than blkCap of them - a dense cluster - keeps no list and
//             re-reads the bins instead).
//   resolve : every thread sweeps the blocker lists of its undecided entries: a rejected blocker is dropped, an
//             accepted one adds ceil(LUT * 0.99 nsc1) and is dropped, an undecided one stays. Then the entry is
//             rejected if nsc1 < min(255, sum) (early: more stamps cannot lower the sum), accepted when the list is
//             empty. States only ever go 0 -> 1 or 0 -> 2 and are the only shared mutable data, so the warps sweep
//             without any barrier until their entries are decided; the lowest undecided entry always has no
//             undecided blocker, hence progress. The number of sweeps is the dependency depth (20-40,
//             tools/select_worklist_model.py) instead of 400-2 500 sequential accept steps with a 961-cell stamp each.
//   The cap (kept == maxNumKpt, uniformity-enforcement-inl.h:199) cuts the accepted list in candidate order after a
//   window: decisions never depend on higher-index candidates, so the first maxKpt accepted are the reference's.
// ----------------------------------------------------------------------------------------------
constexpr int kWlThreads = 1024;
constexpr int kWlCap = kWlThreads;                // window capacity: one entry per thread
constexpr int kWlMaxBins = 2048;                  // padded bin grid (one empty bin all around)
constexpr int kWlMaxBlk = 48, kWlMinBlk = 8;      // blocker list capacity per entry (sized to the free shared memory)

constexpr int kWlMinLut = 16 * 81;                // [py][px][by + 4][bx + 4], see the kernel

struct WlSmem {                                   // dynamic shared memory: 2 x maxOut nodes | WlSmem | lower-bound image | blocker lists
  float lut[31 * 31 + 3];
  unsigned cyx[kWlCap];                           // cy << 16 | cx, window entries in candidate order
  float nsc[kWlCap];                              // nsc1
  int sum[kWlCap];                                // stamps received from the points of earlier windows
  int state[kWlCap];                              // 0 undecided, 1 rejected, 2 accepted
  int idx[kWlCap];                                // candidate index
  int binStart[kWlMaxBins + 4];                   // CSR of the window entries by bin
  int binFill[kWlMaxBins];
  uint2 binEnt[kWlCap];                           // CSR payload: {cy << 16 | cx, position}
  int binStartA[kWlMaxBins + 4];                  // CSR of the points accepted in earlier windows by bin
  unsigned short nblk[kWlCap];                    // blockers found per entry (may exceed the list capacity)
  float minLut[kWlMinLut];                        // smallest LUT value a point at (py, px) of its 4x4 block stamps into block (by, bx)
  unsigned char nzBlock[16][81];                  // per (py, px): the blocks with a positive minLut, as (by + 4) << 4 | (bx + 4)
  int nzCount[16];
  int warpcnt[32], warpoff[32], total, posNext;
};

__device__ __forceinline__ int wl_stamp(const float* lut, int dy, int dx, float nsc) {
  const float v = ceilf(lut[dy * 31 + dx] * nsc);
  return (v == v) ? ((int)v & 0xff) : 0;          // NaN adds 0 (cvttss2si + byte truncation in the reference)
}

#ifdef OKC_SELECT_TRACE
#define WL_T(i) do { if (tid == 0) tr[i] += clock64() - t0; t0 = clock64(); } while (0)
#else
#define WL_T(i) do { } while (0)
#endif

template <int kRegs>   // (see order_candidates_kernel)
__global__ void __maxnreg__(kRegs)
select_keypoints_wave_kernel(const Slot* __restrict__ slots, const float* __restrict__ lut_global, int W, int H,
                             float scaling, unsigned long long maxKpt, int maxOut, int binShift, int nbxP, int nbinsP,
                             int blkCap, int lbW, int lbH) {
  const Slot& slot = slots[blockIdx.x];
  uint64_t* __restrict__ cand = slot.cand;
  int* __restrict__ counters = slot.counters;
  extern __shared__ __align__(16) uint8_t smem[];
  uint2* node = reinterpret_cast<uint2*>(smem);   // accepted points {cy << 16 | cx, 0.99f * nsc1} in acceptance order
  uint2* nodeS = node + maxOut;                   // the same, grouped by bin (CSR binStartA), rebuilt after a window
  WlSmem& S = *reinterpret_cast<WlSmem*>(smem + 2 * (size_t)maxOut * sizeof(uint2));
  // lower bound of the occupancy per 4x4 block of cells (four blocks of margin all around), see step 5
  int* lb = reinterpret_cast<int*>(smem + 2 * (size_t)maxOut * sizeof(uint2) + sizeof(WlSmem));
  // blocker lists, entry-minor: blk[i * kWlCap + e] = i-th remaining blocker of entry e
  unsigned short* blk = reinterpret_cast<unsigned short*>(lb + (size_t)lbW * lbH);
  volatile int* vstate = S.state;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef OKC_SELECT_TRACE
  long long tr[12] = {}, t0 = clock64();
  int trStages = 0;
#endif
  int n = counters[0];
  if (n > kCandCap) n = kCandCap;
  if (counters[3] == 1) n = 0;   // candidate overflow: cand[] was not written (reported as OKC_ERR_CAPACITY)
  if (n == 0) {
    if (tid == 0) counters[1] = 0;
    return;
  }
  for (int i = tid; i < 31 * 31; i += kWlThreads) S.lut[i] = lut_global[i];
  __syncthreads();

  const float maxScore = (float)cand_score(cand[0]);
  const unsigned long long cap = maxKpt ? maxKpt : ~0ull;   // "kept == maxNumKpt" never fires for 0
  const unsigned ltMask = (1u << lane) - 1u;
  unsigned long long kept = 0;
  int pos = 0, sweeps = 0, windows = 0;
  bool overflow = false, lbReady = false;
  // (a positive maximum keeps every stamp value below 256: no byte wrap, which the lower bound could not follow)
  const bool lbUsable = maxScore > 0.0f;

  // exclusive block scan of a flag in thread order; every thread gets the total
  auto scan = [&](bool flag, int& total) -> int {
    const unsigned ballot = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) S.warpcnt[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
      const int c = S.warpcnt[lane];
      int incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      S.warpoff[lane] = incl - c;
      if (lane == 31) S.total = incl;
    }
    __syncthreads();
    const int r = S.warpoff[warp] + __popc(ballot & ltMask);
    total = S.total;
    __syncthreads();   // warpcnt / total may be rewritten by the next scan
    return r;
  };
  auto bin_of = [&](unsigned cyx) -> int {
    return ((int)((cyx >> 16) >> binShift) + 1) * nbxP + ((int)((cyx & 0xffffu) >> binShift) + 1);
  };
  // exclusive scan of S.binFill[0 .. nbinsP) -> start[0 .. nbinsP], two bins per thread (start[nbinsP] = total)
  auto scan_bins = [&](int* start) {
    const int b0 = 2 * tid, b1 = 2 * tid + 1;
    const int c0 = b0 < nbinsP ? S.binFill[b0] : 0, c1 = b1 < nbinsP ? S.binFill[b1] : 0;
    int incl = c0 + c1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) S.warpcnt[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const int c = S.warpcnt[lane];
      int iv = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, iv, d);
        if (lane >= d) iv += t;
      }
      S.warpoff[lane] = iv - c;
    }
    __syncthreads();
    const int ex = S.warpoff[warp] + incl - (c0 + c1);
    if (b0 <= nbinsP) start[b0] = ex;
    if (b1 <= nbinsP) start[b1] = ex + c0;
    __syncthreads();
  };

  while (pos < n && kept < cap && !overflow) {
    // ---- 1. fill the window with candidates in order.
    //   stage  : chunks of 1024 candidates are pre-tested against the lower bound of the occupancy (step 5, one read
    //            each) and the survivors collected until 1024 are staged;
    //   gather : every staged candidate sums the stamps of the points accepted in earlier windows that lie in the 3x3
    //            bins around it (walked as one range, four points in flight) and is dropped as soon as the sum
    //            rejects it: the sum only grows, so the reference would reject it at its turn;
    //   the rest enters the window. ----
    int wcount = 0;
    int* stIdx = reinterpret_cast<int*>(blk);   // staging (the blocker lists are idle here): candidate index,
    uint2* stEnt = S.binEnt;                    // {cell, nsc1}
    while (pos < n && wcount < kWlCap) {
      int staged = 0;
      while (pos < n && staged < kWlCap) {
        const int i = pos + tid;
        bool pass = false;
        unsigned cyx = 0;
        float nsc1 = 0.0f;
        if (i < n) {
          const uint64_t key = cand[i];
          const int score = cand_score(key);
          const uint32_t raster = (uint32_t)key;
          const int y = raster / W, x = raster - y * W;
          const int cy = (int)((float)y * scaling + 16.0f);
          const int cx = (int)((float)x * scaling + 16.0f);
          cyx = ((unsigned)cy << 16) | (unsigned)cx;
          nsc1 = sqrtf(sqrtf((float)score / maxScore)) * 255.0f;
          // the occupancy at the candidate's cell is at least the lower bound of its 4x4 block: a candidate that fails
          // against the bound fails against the occupancy
          pass = !(lbReady && lbUsable && nsc1 < (float)min(lb[((cy >> 2) + 4) * lbW + (cx >> 2) + 4], 255));
        }
        int total;
        const int sp = staged + scan(pass, total);
        if (pass) {
          if (sp < kWlCap) {
            stEnt[sp] = make_uint2(cyx, __float_as_uint(nsc1));
            stIdx[sp] = i;
          } else if (sp == kWlCap) {
            S.posNext = i;   // the stage is full: this chunk's tail goes back to the list
          }
        }
        staged += total;
        pos += kWlThreads;
        if (!lbReady) break;   // (no accepted points yet: nothing to pre-test or gather, a chunk is a stage)
      }
      __syncthreads();
      WL_T(10);
      if (staged > kWlCap) {
        staged = kWlCap;
        pos = S.posNext;
      }
      bool alive = tid < staged;
      unsigned cyx = 0;
      float nsc1 = 0.0f;
      int sum = 0, idx = 0;
      if (alive) {
        const uint2 c = stEnt[tid];
        cyx = c.x;
        nsc1 = __uint_as_float(c.y);
        idx = stIdx[tid];
        if (kept > 0) {
          const int cx = (int)(cyx & 0xffffu), cy = (int)(cyx >> 16);
          const int bin = bin_of(cyx);
          const int a0 = S.binStartA[bin - nbxP - 1], n0 = S.binStartA[bin - nbxP + 2] - a0;
          const int a1 = S.binStartA[bin - 1], n1 = S.binStartA[bin + 2] - a1;
          const int a2 = S.binStartA[bin + nbxP - 1], n2 = S.binStartA[bin + nbxP + 2] - a2;
          const int n01 = n0 + n1, nAll = n01 + n2;
          for (int t = 0; t < nAll && alive; t += 4) {
            uint2 nd[4];   // four points in flight: the loop is a chain of shared-memory latencies otherwise
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int tt = t + u;
              const int j = tt < n0 ? a0 + tt : (tt < n01 ? a1 + (tt - n0) : a2 + (tt - n01));
              nd[u] = tt < nAll ? nodeS[j] : make_uint2(0xffffffffu, 0u);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int dx = cx - (int)(nd[u].x & 0xffffu) + 15, dy = cy - (int)(nd[u].x >> 16) + 15;
              if ((unsigned)dx < 31u && (unsigned)dy < 31u) sum += wl_stamp(S.lut, dy, dx, __uint_as_float(nd[u].y));
            }
            if (nsc1 < (float)min(sum, 255)) alive = false;
          }
        }
      }
      __syncthreads();   // (every thread has read its staged candidate: the stage may be refilled)
      WL_T(11);
#ifdef OKC_SELECT_TRACE
      if (tid == 0) ++trStages;
#endif
      int total;
      const int p = wcount + scan(alive, total);
      if (alive) {
        if (p < kWlCap) {
          S.cyx[p] = cyx;
          S.nsc[p] = nsc1;
          S.sum[p] = sum;
          S.state[p] = 0;
          S.idx[p] = idx;
        } else if (p == kWlCap) {
          S.posNext = idx;   // the window is full: the rest of the stage goes back to the list
        }
      }
      wcount += total;
      if (wcount > kWlCap) {
        __syncthreads();
        pos = S.posNext;
      }
    }
    if (wcount == 0) break;   // pos >= n and nothing left to decide
    for (int b = tid; b < nbinsP; b += kWlThreads) S.binFill[b] = 0;
    ++windows;
    if (wcount > kWlCap) wcount = kWlCap;
    __syncthreads();
    WL_T(1);

    // ---- 2. CSR of the window entries by bin ----
    const int e = tid;           // this thread's entry (position in the window)
    const bool live = e < wcount;
    const unsigned myCyx = live ? S.cyx[e] : 0u;
    const float myNsc = live ? S.nsc[e] : 0.0f;
    const int myBin = bin_of(myCyx);
    const int cx = (int)(myCyx & 0xffffu), cy = (int)(myCyx >> 16);
    const int sum0 = live ? S.sum[e] : 0;
    const int mySlot = live ? atomicAdd(&S.binFill[myBin], 1) : 0;
    __syncthreads();
    scan_bins(S.binStart);
    if (live) S.binEnt[S.binStart[myBin] + mySlot] = make_uint2(myCyx, (unsigned)e);
    __syncthreads();
    WL_T(2);

    // ---- 3. blocker lists: lower-position entries within +-15 cells. Built in CSR order - the lanes of a warp take
    //         entries of the same or the next bin, so they walk (nearly) the same ranges without divergence ----
    if (tid < wcount) {
      const uint2 me = S.binEnt[tid];
      const int ee = (int)me.y;
      const int cxm = (int)(me.x & 0xffffu) - 15, cym = (int)(me.x >> 16) - 15;
      const int bin = bin_of(me.x);
      unsigned short* list = blk + ee;
      int c = 0;
      for (int row = -1; row <= 1; ++row) {
        const int bb = bin + row * nbxP;
        const int j1 = S.binStart[bb + 2];
        for (int j = S.binStart[bb - 1]; j < j1; j += 4) {
          uint2 ent[4];   // four entries in flight
#pragma unroll
          for (int u = 0; u < 4; ++u) ent[u] = j + u < j1 ? S.binEnt[j + u] : make_uint2(0u, 0x7fffffffu);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const unsigned dx = (unsigned)((int)(ent[u].x & 0xffffu) - cxm), dy = (unsigned)((int)(ent[u].x >> 16) - cym);
            if ((int)ent[u].y < ee && dx < 31u && dy < 31u) {
              if (c < blkCap) list[c * kWlCap] = (unsigned short)ent[u].y;
              ++c;
            }
          }
        }
      }
      S.nblk[ee] = (unsigned short)c;
    }
    __syncthreads();
    WL_T(3);

    // ---- 4. resolve: barrier-free sweeps over the blocker lists ----
    {
      bool und = live;
      int nb = live ? (int)S.nblk[e] : 0;   // > blkCap: the list does not hold every blocker (a dense cluster)
      int sum = sum0;
      const float nsc1 = myNsc;
      for (;;) {
        ++sweeps;
        if (und) {
          int total = sum;
          bool blocked;
          if (nb <= blkCap) {
            int w = 0;
            for (int i0 = 0; i0 < nb; i0 += 8) {
              int q[8], st[8];   // eight blockers in flight (two dependent shared-memory reads each)
#pragma unroll
              for (int u = 0; u < 8; ++u) q[u] = i0 + u < nb ? (int)blk[(i0 + u) * kWlCap + e] : 0;
#pragma unroll
              for (int u = 0; u < 8; ++u) st[u] = i0 + u < nb ? vstate[q[u]] : 1;
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                if (st[u] == 0) {
                  blk[w * kWlCap + e] = (unsigned short)q[u];
                  ++w;
                } else if (st[u] == 2) {
                  const unsigned o = S.cyx[q[u]];
                  total += wl_stamp(S.lut, cy - (int)(o >> 16) + 15, cx - (int)(o & 0xffffu) + 15, 0.99f * S.nsc[q[u]]);
                }
              }
            }
            nb = w;
            sum = total;
            blocked = w != 0;
          } else {
            // no complete list: the sum of the blockers accepted so far, from the bins; it is a lower bound of the
            // final sum, and the final sum once no blocker is undecided
            blocked = false;
            for (int row = -1; row <= 1; ++row) {
              const int bb = myBin + row * nbxP;
              const int j1 = S.binStart[bb + 2];
              for (int j = S.binStart[bb - 1]; j < j1; ++j) {
                const uint2 ent = S.binEnt[j];
                const int dx = cx - (int)(ent.x & 0xffffu) + 15, dy = cy - (int)(ent.x >> 16) + 15;
                if ((int)ent.y < e && (unsigned)dx < 31u && (unsigned)dy < 31u) {
                  const int st = vstate[ent.y];
                  if (st == 0) blocked = true;
                  else if (st == 2) total += wl_stamp(S.lut, dy, dx, 0.99f * S.nsc[ent.y]);
                }
              }
            }
          }
          if (nsc1 < (float)min(total, 255)) {
            vstate[e] = 1;
            und = false;
          } else if (!blocked) {
            vstate[e] = 2;
            und = false;
          }
        }
        if (!__any_sync(0xffffffffu, und)) break;
      }
    }
    WL_T(4);
    __syncthreads();
    WL_T(5);

    // ---- 5. the window's accepted entries, in candidate order: kept list (behind the keys) + nodes ----
    const bool acc = live && S.state[e] == 2;
    int totalAcc;
    const int rank = scan(acc, totalAcc);
    if (acc) {
      const unsigned long long k = kept + (unsigned long long)rank;
      if (k < cap && k < (unsigned long long)maxOut) {
        cand[kCandCap + k] = cand[S.idx[e]];
        node[k] = make_uint2(myCyx, __float_as_uint(0.99f * myNsc));
      }
    }
    kept += (unsigned long long)totalAcc;
    if (kept > cap) kept = cap;
    if (kept > (unsigned long long)maxOut) overflow = true;   // flagged below; the host raises OKC_ERR_CAPACITY
    for (int b = tid; b < nbinsP; b += kWlThreads) S.binFill[b] = 0;
    __syncthreads();
    WL_T(6);
    if (pos < n && kept < cap && !overflow) {
      // Another window follows. (a) Lower bound of the occupancy per 4x4 block of cells, the pre-test of the fill: an
      // accepted point at (py, px) inside its block adds at least ceil(minLut[py][px][by][bx] * 0.99 nsc1) to every
      // cell of the block (by, bx) blocks away (ceil is monotone; minLut = the smallest LUT value over the 16 cells).
      if (!lbReady) {
        lbReady = true;
        for (int t = tid; t < kWlMinLut; t += kWlThreads) {
          const int bx = t % 9 - 4, by = (t / 9) % 9 - 4, px = (t / 81) & 3, py = t / 324;
          float m = 1.0f;
          for (int iy = 0; iy < 4; ++iy)
            for (int ix = 0; ix < 4; ++ix) {
              const int dy = 4 * by + iy - py + 15, dx = 4 * bx + ix - px + 15;
              m = fminf(m, ((unsigned)dy < 31u && (unsigned)dx < 31u) ? S.lut[dy * 31 + dx] : 0.0f);
            }
          S.minLut[t] = m;
        }
        for (int t = tid; t < lbW * lbH; t += kWlThreads) lb[t] = 0;
        __syncthreads();
        if (warp < 16) {   // compact list of the blocks a point at (py, px) = warp reaches at all (about 20 of 81)
          int base = 0;
          for (int t0 = 0; t0 < 81; t0 += 32) {
            const int t = t0 + lane;
            const bool nz = t < 81 && S.minLut[warp * 81 + t] > 0.0f;
            const unsigned b = __ballot_sync(0xffffffffu, nz);
            if (nz) S.nzBlock[warp][base + __popc(b & ltMask)] = (unsigned char)(((t / 9) << 4) | (t % 9));
            base += __popc(b);
          }
          if (lane == 0) S.nzCount[warp] = base;
        }
        __syncthreads();
      }
      WL_T(8);
      if (acc) {   // (one thread per accepted entry: about 20 shared-memory atomics each, every lane of a warp busy)
        const int variant = (cy & 3) * 4 + (cx & 3);
        const float* ml = S.minLut + variant * 81;
        int* dst = lb + (cy >> 2) * lbW + (cx >> 2);   // block (by, bx) = (-4, -4) of the margin-shifted image
        const float nsc = 0.99f * myNsc;
        const int nz = S.nzCount[variant];
        for (int t = 0; t < nz; ++t) {
          const int rc = (int)S.nzBlock[variant][t], row = rc >> 4, col = rc & 15;
          const float v = ceilf(ml[row * 9 + col] * nsc);
          const int add = (v == v) ? ((int)v & 0xff) : 0;
          if (add) atomicAdd(dst + row * lbW + col, add);
        }
      }
      WL_T(9);
      // (b) every accepted point regrouped by bin, for the exact sums of the next window's entries
      const int nk = (int)kept;
      for (int k = tid; k < nk; k += kWlThreads) atomicAdd(&S.binFill[bin_of(node[k].x)], 1);
      __syncthreads();
      scan_bins(S.binStartA);
      for (int b = tid; b < nbinsP; b += kWlThreads) S.binFill[b] = S.binStartA[b];   // cursors
      __syncthreads();
      for (int k = tid; k < nk; k += kWlThreads) {
        const uint2 nd = node[k];
        nodeS[atomicAdd(&S.binFill[bin_of(nd.x)], 1)] = nd;
      }
      __syncthreads();
      // (the cursors ended at the bin ends: clear them again for the next window's CSR)
      for (int b = tid; b < nbinsP; b += kWlThreads) S.binFill[b] = 0;
      __syncthreads();
      WL_T(0);
    }
  }

  const int nOut = kept < (unsigned long long)maxOut ? (int)kept : maxOut;
  if (tid == 0) {
    counters[1] = nOut;     // the sub-pixel kernel that follows turns the kept keys into keypoints
    counters[5] = sweeps;   // diagnostic (okc_debug_batch_counter): sweeps of warp 0
    if (kept > (unsigned long long)maxOut) counters[3] = 2;
  }
#ifdef OKC_SELECT_TRACE
  __syncthreads();
  WL_T(7);
  if (tid == 0)
    printf("select trace (cycles): fill %lld csr %lld lists %lld resolve(warp0) %lld wait %lld end %lld regroup %lld "
           "(init %lld stamp %lld) tail %lld sweeps %d windows %d | stages %d: pre-test %lld gather %lld\n", tr[1], tr[2], tr[3], tr[4], tr[5], tr[6], tr[0], tr[8],
           tr[9], tr[7], sweeps, windows, trStages, tr[10], tr[11]);
#endif
}

// Sub-pixel refinement of the kept candidates -> cv::KeyPoint (scale-space-layer-inl.h:394-420), after the selection
// kernel: FP64 work for 700-2 500 points that spreads over the GPU instead of running on the selection CTA's one SM.
__global__ void __launch_bounds__(128)
subpixel_kernel(const Slot* __restrict__ slots, int W, int H, double kpScale, double kpOffset, int octave) {
  const Slot& slot = slots[blockIdx.y];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= slot.counters[1]) return;
  const uint64_t key = slot.cand[kCandCap + i];
  const int score = cand_score(key);
  const uint32_t raster = (uint32_t)key;
  const int v = raster / W, u = raster - v * W;
  int c[9];
  harris_ring_scores(slot.image, W, H, u, v, c);
  float dx, dy;
  subpixel2d((double)c[0], (double)c[1], (double)c[2], (double)c[3], (double)c[4], (double)c[5], (double)c[6],
             (double)c[7], (double)c[8], dx, dy);
  okc_keypoint kp;
  // _scale * ((it->x + delta_x) + _offset): (float + float) widened, the rest in double, narrowed on assignment
  kp.x = (float)(kpScale * ((double)((float)u + dx) + kpOffset));
  kp.y = (float)(kpScale * ((double)((float)v + dy) + kpOffset));
  kp.size = (float)(kpScale * 12.0);
  kp.angle = -1.0f;
  kp.response = (float)score;
  kp.octave = octave;
  kp.class_id = -1;
  slot.kp_detect[i] = kp;
}

// Bytes of per-slot global scratch (Slot::occupancy) the selection kernels need: the occupancy image of the
// sequential-visit kernel when it does not fit in shared memory.
size_t select_scratch_bytes(const DetectArgs& a) {
  const size_t rows = (size_t)((float)(a.height - 3) * a.scaling + 16.0f) + 17;
  const size_t cols = (size_t)((float)(a.width - 3) * a.scaling + 16.0f) + 17;
  const size_t stride = (cols + 15) & ~(size_t)15;
  return rows * stride + 64;
}

void launch_select_keypoints(const Slot* slots, int n, const float* lut, const DetectArgs& a, cudaStream_t s) {
  int dev = 0;
  cudaGetDevice(&dev);
  static int s_dyn[2][64] = {};   // per device: dynamic shared memory granted to the two kernels
  const size_t occRows = (size_t)((float)(a.height - 3) * a.scaling + 16.0f) + 17;
  const size_t occCols = (size_t)((float)(a.width - 3) * a.scaling + 16.0f) + 17;
  for (int which = 0; which < 2; ++which) {
    int& granted = s_dyn[which][dev & 63];
    if (granted != 0) continue;
    int maxOptin = 0;
    cudaDeviceGetAttribute(&maxOptin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa;
    if (which == 1) cudaFuncGetAttributes(&fa, select_keypoints_kernel<true>);
    else cudaFuncGetAttributes(&fa, select_keypoints_wave_kernel<64>);
    granted = (maxOptin - (int)fa.sharedSizeBytes - 1024) & ~15;  // dynamic part beside the static arrays
    if (which == 1)
      cudaFuncSetAttribute(select_keypoints_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, granted);
    else
    {
      cudaFuncSetAttribute(select_keypoints_wave_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, granted);
      cudaFuncSetAttribute(select_keypoints_wave_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, granted);
    }
  }
  // worklist kernel: bins of 2^shift >= 16 occupancy cells, one empty bin all around, at most kWlMaxBins in total
  int shift = 4, nbxP = 0, nbinsP = 0;
  for (;; ++shift) {
    nbxP = (int)((occCols + ((size_t)1 << shift) - 1) >> shift) + 2;
    const int nbyP = (int)((occRows + ((size_t)1 << shift) - 1) >> shift) + 2;
    nbinsP = nbxP * nbyP;
    if (nbinsP <= kWlMaxBins) break;
  }
  // blocker lists take what the accepted-point nodes and the fixed arrays leave of the shared memory
  const int lbW = (int)((occCols + 3) / 4) + 9, lbH = (int)((occRows + 3) / 4) + 9;   // 4x4 blocks, 4 of margin
  const size_t wlFixed = 2 * (size_t)a.max_out * sizeof(uint2) + sizeof(WlSmem) + (size_t)lbW * lbH * sizeof(int);
  int blkCap = 0;
  if (wlFixed < (size_t)s_dyn[0][dev & 63]) {
    blkCap = (int)(((size_t)s_dyn[0][dev & 63] - wlFixed) / ((size_t)kWlCap * sizeof(unsigned short)));
    if (blkCap > kWlMaxBlk) blkCap = kWlMaxBlk;
  }
  // g_knobs.select_sequential: the sequential-visit kernel as the A/B baseline of the tests; it also serves handles
  // whose max_keypoints leave no room for the blocker lists (> ~6 000 keypoints per image)
  if (!g_knobs.select_sequential && blkCap >= kWlMinBlk && occCols < 65536 && occRows < 65536) {
    const size_t wlBytes = wlFixed + (size_t)blkCap * kWlCap * sizeof(unsigned short);
    if (n >= kShareSmMinSlots) select_keypoints_wave_kernel<48><<<n, kWlThreads, wlBytes, s>>>(slots, lut, a.width, a.height, a.scaling, a.max_kpt,
                                                               a.max_out, shift, nbxP, nbinsP, blkCap, lbW, lbH);
    else select_keypoints_wave_kernel<64><<<n, kWlThreads, wlBytes, s>>>(slots, lut, a.width, a.height, a.scaling, a.max_kpt,
                                                               a.max_out, shift, nbxP, nbinsP, blkCap, lbW, lbH);
    subpixel_kernel<<<dim3((a.max_out + 127) / 128, n), 128, 0, s>>>(slots, a.width, a.height, a.kp_scale,
                                                                     a.kp_offset, a.octave);
    return;
  }
  // only as much dynamic shared memory as the occupancy image needs; the global fallback when it does not fit
  const size_t occStride = (occCols + 15) & ~(size_t)15;
  const size_t want = (occRows * occStride + 15) & ~(size_t)15;
  const int smem = want <= (size_t)s_dyn[1][dev & 63] ? (int)want : 0;
  if (smem > 0)
    select_keypoints_kernel<true><<<n, kSelThreads, smem, s>>>(slots, lut, a.width, a.height, a.scaling, a.max_kpt,
                                                               a.max_out, smem, a.kp_scale, a.kp_offset, a.octave);
  else
    select_keypoints_kernel<false><<<n, kSelThreads, 0, s>>>(slots, lut, a.width, a.height, a.scaling, a.max_kpt,
                                                             a.max_out, 0, a.kp_scale, a.kp_offset, a.octave);
}

// ----------------------------------------------------------------------------------------------
// Scale space (octaves > 0) - cold in every shipped config, so built for exactness
// and reuse of the single-layer kernels rather than for speed: one launch chain per layer.
// ----------------------------------------------------------------------------------------------
// brisk::Halfsample8 (brisk/src/image-down-sampling.cc:161-415, SSE path): three rounding rules by column region.
//   body     (dst col < 16*(cols/16/2)): avg(avg_v(2c), avg_v(2c+1)), avg(a,b) = (a+b+1)>>1, avg_v = vertical avg
//   half end (cols/16 odd, next 8 dst cols): (avg_v(2c) + avg_v(2c+1)) / 2   (truncating)
//   leftover ((cols%16)/2 dst cols): (a + b + c + d + 2) / 4
__global__ void halfsample8_kernel(const uint8_t* __restrict__ src, int sw, int sh, uint8_t* __restrict__ dst) {
  const int dw = sw / 2, dh = sh / 2;
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c >= dw || r >= dh) return;
  const int hsize = sw / 16, bodyCols = 16 * (hsize / 2), halfCols = (hsize & 1) ? 8 : 0;
  const uint8_t* p1 = src + (size_t)(2 * r) * sw + 2 * c;
  const uint8_t* p2 = p1 + sw;
  const int a = p1[0], b = p1[1], cc = p2[0], d = p2[1];
  int v;
  if (c < bodyCols) {
    const int v0 = (a + cc + 1) >> 1, v1 = (b + d + 1) >> 1;
    v = (v0 + v1 + 1) >> 1;
  } else if (c < bodyCols + halfCols) {
    const int v0 = (a + cc + 1) >> 1, v1 = (b + d + 1) >> 1;
    v = (v0 + v1) / 2;
  } else {
    v = (a + b + cc + d + 2) / 4;
  }
  dst[(size_t)r * dw + c] = (uint8_t)v;
}

// brisk::Twothirdsample8 (:573-814, SSE path). Per 3x3 source block -> 2x2 destination block.
//   body (src col block < cols/15): vertical u = avg(avg(first, second), first) (lower: third), horizontal
//        out[2t] = avg(avg(u[3t], u[3t+1]), u[3t]), out[2t+1] = avg(avg(u[3t+2], u[3t+1]), u[3t+2])
//   remainder: (4*A1 + 2*(A2 + B1 + 1) + B2 + 1) / 9 and its mirrored forms
__global__ void twothirdsample8_kernel(const uint8_t* __restrict__ src, int sw, int sh, uint8_t* __restrict__ dst) {
  const int dw = (sw / 3) * 2, dh = (sh / 3) * 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x, rb = blockIdx.y;  // 3x3 block column / row
  if (t >= sw / 3 || rb >= sh / 3) return;
  const uint8_t* p = src + (size_t)(3 * rb) * sw + 3 * t;
  const int A1 = p[0], A2 = p[1], A3 = p[2], B1 = p[sw], B2 = p[sw + 1], B3 = p[sw + 2];
  const int C1 = p[2 * sw], C2 = p[2 * sw + 1], C3 = p[2 * sw + 2];
  int o00, o01, o10, o11;
  if (3 * t < 15 * (sw / 15)) {
    auto avg = [](int x, int y) { return (x + y + 1) >> 1; };
    const int u0 = avg(avg(A1, B1), A1), u1 = avg(avg(A2, B2), A2), u2 = avg(avg(A3, B3), A3);
    const int l0 = avg(avg(C1, B1), C1), l1 = avg(avg(C2, B2), C2), l2 = avg(avg(C3, B3), C3);
    o00 = avg(avg(u0, u1), u0);
    o01 = avg(avg(u2, u1), u2);
    o10 = avg(avg(l0, l1), l0);
    o11 = avg(avg(l2, l1), l2);
  } else {
    o00 = ((4 * A1 + 2 * (A2 + B1 + 1) + B2 + 1) / 9) & 0xff;
    o01 = ((4 * A3 + 2 * (A2 + B3 + 1) + B2 + 1) / 9) & 0xff;
    o10 = ((4 * C1 + 2 * (C2 + B1 + 1) + B2 + 1) / 9) & 0xff;
    o11 = ((4 * C3 + 2 * (C2 + B3 + 1) + B2 + 1) / 9) & 0xff;
  }
  uint8_t* q = dst + (size_t)(2 * rb) * dw + 2 * t;
  q[0] = (uint8_t)o00;
  q[1] = (uint8_t)o01;
  q[dw] = (uint8_t)o10;
  q[dw + 1] = (uint8_t)o11;
}

void launch_halfsample8(const uint8_t* src, int sw, int sh, uint8_t* dst, cudaStream_t s) {
  if (sw / 2 <= 0 || sh / 2 <= 0) return;
  dim3 grid((sw / 2 + 127) / 128, sh / 2);
  halfsample8_kernel<<<grid, 128, 0, s>>>(src, sw, sh, dst);
}
void launch_twothirdsample8(const uint8_t* src, int sw, int sh, uint8_t* dst, cudaStream_t s) {
  if (sw / 3 <= 0 || sh / 3 <= 0) return;
  dim3 grid((sw / 3 + 127) / 128, sh / 3);
  twothirdsample8_kernel<<<grid, 128, 0, s>>>(src, sw, sh, dst);
}

// HarrisScoreCalculator::Score(double u, double v) (brisk/include/brisk/harris-score-calculator.h:69-87)
__device__ __forceinline__ double score_bilinear(const int32_t* __restrict__ sc, int W, int H, double u, double v) {
  const int ui = (int)u, vi = (int)v;
  if (ui + 1 >= W || vi + 1 >= H || ui < 0 || vi < 0) return 0.0;
  const double ru = u - (double)ui, rv = v - (double)vi;
  const double omu = 1.0 - ru, omv = 1.0 - rv;
  const int32_t* p = sc + (size_t)vi * W + ui;
  return omv * (omu * (double)p[0] + ru * (double)p[1]) + rv * (omu * (double)p[W] + ru * (double)p[W + 1]);
}

// 3-D non-maximum suppression of ScaleSpaceLayer::DetectScaleSpaceMaxima (scale-space-layer-inl.h:230-378): a 2-D
// maximum survives only if it is not smaller than 9 bilinear samples of the layer above and/or below.
// one_over_scale_* is the reference's int truncation of 1.0 / scale (1 above, 0 below - the nine "below" probes
// coincide). Candidates live in the NMS bitmap; failing bits are cleared in place, raster order is untouched.
struct LayerLink {
  const int32_t* scores;
  int W, H;
  double scale, offset;
  int step;
};
__global__ void scale_nms_kernel(const int32_t* __restrict__ scores, uint32_t* __restrict__ bitmap, int W, int H,
                                 LayerLink above, LayerLink below) {
  const int wordsPerRow = harris_strips(W) * kStripWords;
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= H * wordsPerRow) return;
  unsigned bits = bitmap[w];
  if (!bits) return;
  const int y = w / wordsPerRow, rem = w - y * wordsPerRow;
  const int xb = (rem / kStripWords) * kStripCols - 4 + ((rem % kStripWords) << 5);
  unsigned keep = bits;
  while (bits) {
    const int b = __ffs(bits) - 1;
    bits &= bits - 1;
    const int x = xb + b;
    const double center = (double)scores[(size_t)y * W + x];
    bool ok = true;
    const int dxs[9] = {0, 1, -1, 0, 0, 1, 1, -1, -1}, dys[9] = {0, 0, 0, 1, -1, 1, -1, 1, -1};
    if (above.scores) {
      for (int i = 0; i < 9 && ok; ++i) {
        const double u = (double)(x + dxs[i] * above.step), v = (double)(y + dys[i] * above.step);
        if (center < score_bilinear(above.scores, above.W, above.H, above.scale * (u + above.offset),
                                    above.scale * (v + above.offset)))
          ok = false;
      }
    }
    if (ok && below.scores) {
      for (int i = 0; i < 9 && ok; ++i) {
        const double u = (double)(x + dxs[i] * below.step), v = (double)(y + dys[i] * below.step);
        if (center < score_bilinear(below.scores, below.W, below.H, below.scale * (u + below.offset),
                                    below.scale * (v + below.offset)))
          ok = false;
      }
    }
    if (!ok) keep &= ~(1u << b);
  }
  bitmap[w] = keep;
}

void launch_scale_nms(const int32_t* scores, uint32_t* bitmap, int W, int H, const int32_t* aboveScores, int aW,
                      int aH, double aScale, double aOffset, int aStep, const int32_t* belowScores, int bW, int bH,
                      double bScale, double bOffset, int bStep, cudaStream_t s) {
  LayerLink a{aboveScores, aW, aH, aScale, aOffset, aStep}, b{belowScores, bW, bH, bScale, bOffset, bStep};
  const int words = (int)harris_bitmap_words(W, H);
  scale_nms_kernel<<<(words + 127) / 128, 128, 0, s>>>(scores, bitmap, W, H, a, b);
}

// Appends the per-layer keypoints in layer order (detectImpl's loop, scale-space-feature-detector.h:157-162).
__global__ void concat_layers_kernel(const Slot* __restrict__ layers, int nLayers, Slot out, int maxOut) {
  __shared__ int s_off[33];
  if (threadIdx.x == 0) {
    int off = 0, flag = 0;
    for (int l = 0; l < nLayers; ++l) {
      s_off[l] = off;
      off += layers[l].counters[1];
      if (layers[l].counters[3]) flag = layers[l].counters[3];
    }
    s_off[nLayers] = off;
    if (off > maxOut) flag = 2;
    out.counters[0] = layers[0].counters[0];
    out.counters[1] = off < maxOut ? off : maxOut;
    if (flag) out.counters[3] = flag;
  }
  __syncthreads();
  for (int l = 0; l < nLayers; ++l) {
    const int n = s_off[l + 1] - s_off[l];
    for (int i = threadIdx.x; i < n; i += blockDim.x)
      if (s_off[l] + i < maxOut) out.kp_detect[s_off[l] + i] = layers[l].kp_detect[i];
  }
}

void launch_concat_layers(const Slot* layers, int nLayers, const Slot& out, int maxOut, cudaStream_t s) {
  concat_layers_kernel<<<1, 256, 0, s>>>(layers, nLayers, out, maxOut);
}

}  // namespace okc
