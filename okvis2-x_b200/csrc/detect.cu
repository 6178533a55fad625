// Keypoint detection kernels.
//
//  harris_nms_kernel   = brisk::HarrisScoresSSE (brisk/src/harris-scores.cc:65-301) fused with
//                        HarrisScoreCalculator::Get2dMaxima (brisk/src/harris-score-calculator.cc:70-119):
//                        one pass over the u8 image, no int16 planes in HBM; writes the int32 score map
//                        (needed by the sub-pixel fit) and a bitmap of the NMS survivors.
//  order_candidates_kernel = the std::sort at the top of EnforceKeyPointUniformity, tie order included
//                        (libstdc++ introsort reproduced exactly; see the comment at the kernel).
//  select_keypoints_kernel = the greedy loop of EnforceKeyPointUniformity
//                        (brisk/include/brisk/internal/uniformity-enforcement-inl.h:57-207)
//                        + the Subpixel2D tail of DetectScaleSpaceMaxima
//                        (brisk/include/brisk/internal/scale-space-layer-inl.h:394-420, 568-701):
//                        one CTA per image runs the greedy occupancy loop with the occupancy image in
//                        shared memory, then refines.
#include "common.cuh"

namespace okc {

// ----------------------------------------------------------------------------------------------
// Harris score + 8-neighbour NMS
// ----------------------------------------------------------------------------------------------
constexpr int TX = 32;   // score tile width  (outputs per CTA)
constexpr int TY = 16;   // score tile height
constexpr int PXW = TX + 6, PXH = TY + 6;   // pixel tile, halo 3
constexpr int PRW = TX + 4, PRH = TY + 4;   // gradient-product tile, halo 2
constexpr int SCW = TX + 2, SCH = TY + 2;   // score tile, halo 1 (NMS neighbours)

__global__ void __launch_bounds__(256) harris_nms_kernel(const Slot* __restrict__ slots, int W, int H,
                                                         int absThreshold) {
  const Slot& slot = slots[blockIdx.z];
  const uint8_t* __restrict__ img = slot.image;
  int32_t* __restrict__ scores = slot.scores;
  uint32_t* __restrict__ bitmap = slot.bitmap;
  __shared__ uint8_t px[PXH][PXW + 2];
  __shared__ short pxx[PRH][PRW], pxy[PRH][PRW], pyy[PRH][PRW];
  __shared__ int sc[SCH][SCW];

  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;

  // 1. pixel tile (zero outside the image; such pixels never reach a valid score)
  for (int i = tid; i < PXH * PXW; i += 256) {
    int ly = i / PXW, lx = i - ly * PXW;
    int gx = x0 - 3 + lx, gy = y0 - 3 + ly;
    uint8_t v = 0;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) v = __ldg(img + (size_t)gy * W + gx);
    px[ly][lx] = v;
  }
  __syncthreads();

  // 2. Scharr gradients (<<3, int16) and their products (mulhi = >>16), harris-scores.cc:124-206.
  //    Defined on rows 1..H-2, cols 1..W-2 only.
  for (int i = tid; i < PRH * PRW; i += 256) {
    int ly = i / PRW, lx = i - ly * PRW;
    int gx = x0 - 2 + lx, gy = y0 - 2 + ly;
    short vxx = 0, vxy = 0, vyy = 0;
    if (gx >= 1 && gx <= W - 2 && gy >= 1 && gy <= H - 2) {
      int cy = ly + 1, cx = lx + 1;  // centre in px[][]
      int a00 = px[cy - 1][cx - 1], a01 = px[cy - 1][cx], a02 = px[cy - 1][cx + 1];
      int a10 = px[cy][cx - 1], a12 = px[cy][cx + 1];
      int a20 = px[cy + 1][cx - 1], a21 = px[cy + 1][cx], a22 = px[cy + 1][cx + 1];
      int dx = (10 * (a10 - a12) + 3 * (a00 - a02) + 3 * (a20 - a22)) << 3;
      int dy = (10 * (a01 - a21) + 3 * (a00 - a20) + 3 * (a02 - a22)) << 3;
      vxx = (short)((dx * dx) >> 16);
      vyy = (short)((dy * dy) >> 16);
      vxy = (short)((dx * dy) >> 16);
    }
    pxx[ly][lx] = vxx;
    pxy[ly][lx] = vxy;
    pyy[ly][lx] = vyy;
  }
  __syncthreads();

  // 3. 3x3 binomial smoothing (>>4, arithmetic) and the score, harris-scores.cc:231-296.
  //    Valid on rows 2..H-3, cols 2..W-3; the 2-pixel frame stays 0 (Mat::zeros).
  for (int i = tid; i < SCH * SCW; i += 256) {
    int ly = i / SCW, lx = i - ly * SCW;
    int gx = x0 - 1 + lx, gy = y0 - 1 + ly;
    int s = 0;
    if (gx >= 2 && gx <= W - 3 && gy >= 2 && gy <= H - 3) {
      int cy = ly + 1, cx = lx + 1;  // centre in the product tiles
#define OKC_SMOOTH(P)                                                                             \
  ((4 * (int)P[cy][cx] + 2 * ((int)P[cy - 1][cx] + (int)P[cy + 1][cx] + (int)P[cy][cx - 1] + (int)P[cy][cx + 1]) + \
    ((int)P[cy - 1][cx - 1] + (int)P[cy - 1][cx + 1] + (int)P[cy + 1][cx - 1] + (int)P[cy + 1][cx + 1])) >> 4)
      int dxdx = OKC_SMOOTH(pxx);
      int dxdy = OKC_SMOOTH(pxy);
      int dydy = OKC_SMOOTH(pyy);
#undef OKC_SMOOTH
      int t = (dxdx + dydy) >> 1;
      s = dxdx * dydy - dxdy * dxdy - (t * t) / 5;
    }
    sc[ly][lx] = s;
  }
  __syncthreads();

  // 4. write scores; NMS on rows [2, H-2), cols [2, W-2): keep if >= threshold and no neighbour is
  //    strictly greater (plateaus survive), harris-score-calculator.cc:76-117. Survivors are published as a
  //    bitmap (one word per 32-pixel tile row), so that the next kernel can rebuild the reference's RASTER
  //    candidate order without a sort. Each warp owns one tile row per iteration (TX == 32).
  const int wordsPerRow = (W + 31) >> 5;
  for (int i = tid; i < TX * TY; i += 256) {
    int ly = i / TX, lx = i - ly * TX;
    int gx = x0 + lx, gy = y0 + ly;
    bool isMax = false;
    if (gx < W && gy < H) {
      int cy = ly + 1, cx = lx + 1;
      int c = sc[cy][cx];
      scores[(size_t)gy * W + gx] = c;
      if (gx >= 2 && gx < W - 2 && gy >= 2 && gy < H - 2 && c >= absThreshold)
        isMax = !(sc[cy][cx + 1] > c || sc[cy][cx - 1] > c || sc[cy + 1][cx] > c || sc[cy - 1][cx] > c ||
                  sc[cy + 1][cx + 1] > c || sc[cy + 1][cx - 1] > c || sc[cy - 1][cx + 1] > c ||
                  sc[cy - 1][cx - 1] > c);
    }
    const unsigned word = __ballot_sync(0xffffffffu, isMax);
    if (lx == 0 && gy < H) bitmap[(size_t)gy * wordsPerRow + blockIdx.x] = word;
  }
}

void launch_harris_nms(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s) {
  dim3 grid((a.width + TX - 1) / TX, (a.height + TY - 1) / TY, n);
  harris_nms_kernel<<<grid, dim3(32, 8), 0, s>>>(slots, a.width, a.height, a.abs_threshold);
}

__device__ __forceinline__ void bitonic_sort(uint64_t* keys, int npad) {
  // keys[0..npad) ascending; npad is a power of two; all threads of the CTA participate.
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npad; i += 1024) {
        int ixj = i ^ j;
        if (ixj > i) {
          uint64_t a = keys[i], b = keys[ixj];
          bool up = (i & k) == 0;
          if ((a > b) == up) {
            keys[i] = b;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
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
//     array the partition phase leaves behind -> one bitonic sort on (score desc, position asc) keys.
// tools/introsort_model.py checks this formulation against the real std::sort (oracle/_ref) on tie-heavy and
// adversarial (heapsort-triggering) inputs.
constexpr int kOrdThreads = 1024;
constexpr int kOrdSmemCap = 8192;                         // candidates handled entirely in shared memory
constexpr int kOrdQueueSmem = 512;                        // >= kOrdSmemCap / 17 ranges per level
constexpr int kOrdQueueGlobal = kCandCap / 16;            // >= kCandCap / 17

struct OrdArrays {
  int* sc;          // score
  uint32_t* ra;     // raster index y * W + x
  int* lp;          // left stops of the range being partitioned (positions), stored inside the range
  int* rp;          // right stops
  int2* q[2];       // range queues of the current / next level
  uint64_t* keys;   // final sort keys (may alias sc/lp/rp: they are dead by then)
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
__device__ void ord_heap_sort(const OrdArrays& A, int f, int l) {
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
__device__ int ord_warp_partition(const OrdArrays& A, int f, int l, int lane) {
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

__global__ void __launch_bounds__(kOrdThreads, 1)
order_candidates_kernel(const Slot* __restrict__ slots, int W, int H) {
  const Slot& slot = slots[blockIdx.x];
  const int32_t* __restrict__ scores = slot.scores;
  const uint32_t* __restrict__ bitmap = slot.bitmap;
  uint64_t* __restrict__ cand = slot.cand;
  int* __restrict__ counters = slot.counters;
  uint8_t* __restrict__ gscratch = slot.order_scratch;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_warp[32];
  __shared__ int s_n, s_qcount[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- A. candidate count and raster-ordered offsets from the bitmap ----
  const int wordsPerRow = (W + 31) >> 5;
  const int words = H * wordsPerRow;
  const int chunk = (words + kOrdThreads - 1) / kOrdThreads;
  const int w0 = tid * chunk, w1 = min(words, w0 + chunk);
  int mine = 0;
  for (int w = w0; w < w1; ++w) mine += __popc(bitmap[w]);
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

  OrdArrays A;
  const bool inSmem = n <= kOrdSmemCap;
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
    // shared memory: the keys overlay sc/lp/rp (12 * cap bytes >= 8 * cap); global: their own region
    A.keys = inSmem ? reinterpret_cast<uint64_t*>(u) : reinterpret_cast<uint64_t*>(u + cap * 12);
  }
  {
    int pos = s_warp[warp] + incl - mine;
    for (int w = w0; w < w1; ++w) {
      unsigned bits = bitmap[w];
      const int y = w / wordsPerRow, xb = (w - y * wordsPerRow) << 5;
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const uint32_t raster = (uint32_t)(y * W + xb + b);
        A.sc[pos] = scores[raster];
        A.ra[pos] = raster;
        ++pos;
      }
    }
  }
  if (tid == 0) {
    A.q[0][0] = make_int2(0, n);
    s_qcount[0] = n > 16 ? 1 : 0;
    s_qcount[1] = 0;
  }
  __syncthreads();

  // ---- B. introsort partition phase, level by level ----
  int depthLimit = 2 * (31 - __clz(n));
  int cur = 0;
  while (true) {
    const int cnt = s_qcount[cur];
    if (cnt == 0) break;
    for (int s = warp; s < cnt; s += kOrdThreads / 32) {
      const int2 r = A.q[cur][s];
      if (depthLimit == 0) {
        if (lane == 0) ord_heap_sort(A, r.x, r.y);
        continue;
      }
      const int cut = ord_warp_partition(A, r.x, r.y, lane);
      if (lane == 0) {
        if (cut - r.x > 16) A.q[cur ^ 1][atomicAdd(&s_qcount[cur ^ 1], 1)] = make_int2(r.x, cut);
        if (r.y - cut > 16) A.q[cur ^ 1][atomicAdd(&s_qcount[cur ^ 1], 1)] = make_int2(cut, r.y);
      }
    }
    __syncthreads();
    if (tid == 0) s_qcount[cur] = 0;
    cur ^= 1;
    --depthLimit;
    __syncthreads();
  }

  // ---- C. __final_insertion_sort == stable sort by score (desc) of what the partitions left ----
  int npad = 1;
  while (npad < n) npad <<= 1;
  if (inSmem) {
    int mys[kOrdSmemCap / kOrdThreads];
#pragma unroll
    for (int k = 0; k < kOrdSmemCap / kOrdThreads; ++k) {
      const int i = tid + k * kOrdThreads;
      mys[k] = i < n ? A.sc[i] : 0;
    }
    __syncthreads();  // sc/lp/rp are dead from here on: the keys overlay them
#pragma unroll
    for (int k = 0; k < kOrdSmemCap / kOrdThreads; ++k) {
      const int i = tid + k * kOrdThreads;
      if (i < npad) A.keys[i] = i < n ? cand_key(mys[k], (uint32_t)i) : ~0ull;
    }
  } else {
    for (int i = tid; i < npad; i += kOrdThreads) A.keys[i] = i < n ? cand_key(A.sc[i], (uint32_t)i) : ~0ull;
  }
  __syncthreads();
  bitonic_sort(A.keys, npad);
  for (int i = tid; i < n; i += kOrdThreads) {
    const uint64_t k = A.keys[i];
    cand[i] = cand_key(cand_score(k), A.ra[(uint32_t)k]);
  }
}

constexpr size_t kOrdSmemBytes = (size_t)kOrdSmemCap * 4 + 2 * (size_t)kOrdQueueSmem * 8 + (size_t)kOrdSmemCap * 12;
size_t order_scratch_bytes() {
  return (size_t)kCandCap * 4 + 2 * (size_t)kOrdQueueGlobal * 8 + (size_t)kCandCap * 12 + (size_t)kCandCap * 8 + 256;
}

void launch_order_candidates(const Slot* slots, int n, const DetectArgs& a, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(order_candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kOrdSmemBytes);
    configured = true;
  }
  order_candidates_kernel<<<n, kOrdThreads, kOrdSmemBytes, s>>>(slots, a.width, a.height);
}

// ----------------------------------------------------------------------------------------------
// greedy uniformity + sub-pixel refinement: one CTA per image
// ----------------------------------------------------------------------------------------------
constexpr int kSelThreads = 1024;

// Subpixel2D, scale-space-layer-inl.h:568-701. Returns only the deltas (the refined peak value is
// computed and discarded by the caller in the reference, :403-411).
__device__ void subpixel2d(double s_0_0, double s_0_1, double s_0_2, double s_1_0, double s_1_1, double s_1_2,
                           double s_2_0, double s_2_1, double s_2_2, float& delta_x, float& delta_y) {
  double tmp1 = s_0_0 + s_0_2 - 2 * s_1_1 + s_2_0 + s_2_2;
  double coeff1 = 3 * (tmp1 + s_0_1 - ((s_1_0 + s_1_2) / 2.0) + s_2_1);
  double coeff2 = 3 * (tmp1 - ((s_0_1 + s_2_1) / 2.0) + s_1_0 + s_1_2);
  double tmp2 = s_0_2 - s_2_0;
  double tmp3 = (s_0_0 + tmp2 - s_2_2);
  double tmp4 = tmp3 - 2 * tmp2;
  double coeff3 = -3 * (tmp3 + s_0_1 - s_2_1);
  double coeff4 = -3 * (tmp4 + s_1_0 - s_1_2);
  double coeff5 = (s_0_0 - s_0_2 - s_2_0 + s_2_2) / 4.0;
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

__global__ void __launch_bounds__(kSelThreads, 1)
select_keypoints_kernel(const Slot* __restrict__ slots, const float* __restrict__ lut_global, int W, int H,
                        float scaling, unsigned long long maxKpt, int maxOut, int smemBytes) {
  const Slot& slot = slots[blockIdx.x];
  const int32_t* __restrict__ scores = slot.scores;
  uint64_t* __restrict__ cand = slot.cand;
  int* __restrict__ counters = slot.counters;
  uint8_t* __restrict__ occ_global = slot.occupancy;
  okc_keypoint* __restrict__ out = slot.kp_detect;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ float lut[31 * 31];
  __shared__ int s_off[kSelThreads];
  __shared__ float s_nsc[kSelThreads];
  __shared__ int s_idx[kSelThreads];
  __shared__ int s_warpcnt[32], s_warpoff[32], s_total;

  const int tid = threadIdx.x;
  int n = counters[0];
  if (n > kCandCap) n = kCandCap;
  if (tid < 31 * 31) lut[tid] = lut_global[tid];
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
  uint8_t* occ = occBytes <= (size_t)smemBytes ? smem : occ_global;
  for (size_t i = (size_t)tid * 16; i < occBytes; i += (size_t)kSelThreads * 16)
    *reinterpret_cast<uint4*>(occ + i) = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const float maxScore = (float)cand_score(cand[0]);
  const int ly = tid / 31, lx = tid - ly * 31;
  const float myLut = tid < 31 * 31 ? lut[tid] : 0.0f;
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
    if (i < n) {
      const uint64_t key = cand[i];
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
      int c = s_warpcnt[lane], incl = c;
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
      s_idx[pos] = i;
    }
    __syncthreads();
    const int total = s_total;
    for (int j = 0; j < total; ++j) {
      const int o = s_off[j];
      const float v1 = s_nsc[j];
      if (v1 < (float)occ[o]) continue;
      const float nsc = 0.99f * v1;
      __syncthreads();  // everyone has read the occupancy before the stamp changes it
      if (tid < 31 * 31) {
        float v = ceilf(myLut * nsc);
        int add = (v == v) ? ((int)v & 0xff) : 0;  // NaN -> 0, as cvttss2si + byte truncation
        uint8_t* p = occ + o + (ly - 15) * occStride + (lx - 15);
        int sum = (int)*p + add;
        *p = (uint8_t)(sum > 255 ? 255 : sum);
      }
      if (tid == 0 && kept < maxOut) cand[kCandCap + kept] = cand[s_idx[j]];  // kept list lives behind the keys
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
    const int32_t* c = scores + (size_t)v * W + u;
    float dx, dy;
    subpixel2d((double)c[-W - 1], (double)c[-W], (double)c[-W + 1], (double)c[-1], (double)c[0], (double)c[1],
               (double)c[W - 1], (double)c[W], (double)c[W + 1], dx, dy);
    okc_keypoint kp;
    // _scale * ((it->x + delta_x) + _offset): (float)(1.0 * ((double)((float)x + dx) + 0.0))
    kp.x = (float)(1.0 * ((double)((float)u + dx) + 0.0));
    kp.y = (float)(1.0 * ((double)((float)v + dy) + 0.0));
    kp.size = 12.0f;
    kp.angle = -1.0f;
    kp.response = (float)score;
    kp.octave = 0;
    kp.class_id = -1;
    out[i] = kp;
  }
}

static int g_select_smem = 0;

void launch_select_keypoints(const Slot* slots, int n, const float* lut, const DetectArgs& a, cudaStream_t s) {
  if (g_select_smem == 0) {
    int dev = 0, maxOptin = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&maxOptin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, select_keypoints_kernel);
    g_select_smem = (maxOptin - (int)fa.sharedSizeBytes - 1024) & ~15;  // dynamic part beside the static arrays
    cudaFuncSetAttribute(select_keypoints_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, g_select_smem);
  }
  // only as much dynamic shared memory as the occupancy image needs (more CTAs per SM when several images
  // are selected at once); the global fallback is used when it does not fit
  const size_t occRows = (size_t)((float)(a.height - 3) * a.scaling + 16.0f) + 17;
  const size_t occStride = (((size_t)((float)(a.width - 3) * a.scaling + 16.0f) + 17) + 15) & ~(size_t)15;
  size_t want = occRows * occStride;
  int smem = want <= (size_t)g_select_smem ? (int)((want + 15) & ~(size_t)15) : 0;
  select_keypoints_kernel<<<n, kSelThreads, smem, s>>>(slots, lut, a.width, a.height, a.scaling, a.max_kpt, a.max_out,
                                                       smem);
}

}  // namespace okc
