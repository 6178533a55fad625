// Descriptor extraction kernels: one warp per keypoint.
//
// Follows brisk::BriskDescriptorExtractor::doDescriptorComputation
// (brisk/src/brisk-descriptor-extractor.cc:575-805) and SmoothedIntensity<uchar,int> (:355-535).
// Every float/double operation is written in the reference's evaluation order and precision
// (the file is compiled with --fmad=false); int sums wrap like the reference's int.
//
// The reference builds an integral image (brisk/include/brisk/internal/integral-image.h:66-174) and reads
// 12 taps of it per sample when the box is larger than 2 px; the box sums are exact integers, so here each
// lane adds the pixels of its box directly (<= ~12x12 u8 reads, L1/L2 hits) and no (H+1)x(W+1) int32 image
// is ever written to HBM. The weighted sum is identical term by term.
#include "common.cuh"
#include <type_traits>

namespace okc {

constexpr int kWarpsPerBlock = 8;

// ---- border filter: ordered compaction (brisk-descriptor-extractor.cc:601-627, RoiPredicate :537-541) ----
__global__ void __launch_bounds__(1024) border_filter_kernel(const Slot* __restrict__ slots, int W, int H, int border,
                                                             int maxOut) {
  const Slot& slot = slots[blockIdx.x];
  const okc_keypoint* __restrict__ in = slot.kp_detect;
  okc_keypoint* __restrict__ out = slot.kp_desc;
  int* __restrict__ counters = slot.counters;
  __shared__ int s_warpcnt[32], s_warpoff[32], s_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = counters[1];
  const float minX = (float)border, minY = (float)border, maxX = (float)(W - border), maxY = (float)(H - border);
  int base_out = 0;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    bool keep = false;
    okc_keypoint kp;
    if (i < n) {
      kp = in[i];
      keep = !((kp.x < minX) || (kp.x >= maxX) || (kp.y < minY) || (kp.y >= maxY));
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
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
    if (keep) {
      const int pos = base_out + s_warpoff[warp] + __popc(ballot & ((1u << lane) - 1u));
      if (pos < maxOut) out[pos] = kp;
    }
    base_out += s_total;
    __syncthreads();
  }
  if (tid == 0) counters[2] = base_out < maxOut ? base_out : maxOut;
}

// ---- SmoothedIntensity<unsigned char, int> (brisk-descriptor-extractor.cc:355-535) ----
// Split in two so that a box can be summed by one lane or row-wise by several: box_params() does the float -> int
// weight arithmetic once, box_row() returns the integer contribution of one box row. The reference's result is a
// linear combination of pixel sums in int32, so the order of the additions is free.
struct BoxParams {
  int mode;                  // 0: box leaves the image (-1), 1: sigma < 0.5 bilinear tap (value ready), 2: weighted box
  int value;                 // result for modes 0 and 1
  int x_left, y_top, dx, dy;
  int A, B, C, D;            // corner weights
  int wx0, wy0, wx1, wy1;    // r_x_1_i (left column), r_y_1_i (top row), r_x1_i (right column), r_y1_i (bottom row)
  int scaling, scaling2;
};

__device__ __forceinline__ BoxParams box_params(const uint8_t* __restrict__ img, int W, int H, float key_x, float key_y,
                                                PatternPoint p0, bool warped, const float* warp, float warpScale) {
  BoxParams b;
  b.mode = 0;
  b.value = -1;
  float bx = p0.x, by = p0.y, bs = p0.sigma;
  if (warped) {
    bx = warp[0] * p0.x + warp[1] * p0.y;
    by = warp[2] * p0.x + warp[3] * p0.y;
    bs = warpScale * p0.sigma;
  }
  const float xf = bx + key_x;
  const float yf = by + key_y;
  const int x = (int)xf;
  const int y = (int)yf;
  const float sigma_half = bs;
  const float area = (float)(4.0 * (double)sigma_half * (double)sigma_half);
  const float x_1 = xf - sigma_half, x1 = xf + sigma_half;
  const float y_1 = yf - sigma_half, y1 = yf + sigma_half;

  if (sigma_half < 0.5f) {
    if (x < 0 || x > W - 2 || y < 0 || y > H - 2) return b;
    const int r_x = (int)((xf - (float)x) * 1024.0f);
    const int r_y = (int)((yf - (float)y) * 1024.0f);
    const int r_x_1 = 1024 - r_x, r_y_1 = 1024 - r_y;
    const uint8_t* p = img + x + (size_t)y * W;
    int ret = r_x_1 * r_y_1 * (int)p[0];
    ret += r_x * r_y_1 * (int)p[1];
    ret += r_x * r_y * (int)p[W + 1];
    ret += r_x_1 * r_y * (int)p[W];
    b.mode = 1;
    b.value = ret / 1024;
    return b;
  }
  if (x_1 < 0.0f) return b;
  if (x1 > (float)(W - 1)) return b;
  if (y_1 < 0.0f) return b;
  if (y1 > (float)(H - 1)) return b;

  b.mode = 2;
  b.scaling = (int)(4194304.0 / (double)area);
  b.scaling2 = (int)((double)((float)b.scaling * area) / 1024.0);
  b.x_left = (int)((double)x_1 + 0.5);
  b.y_top = (int)((double)y_1 + 0.5);
  const int x_right = (int)((double)x1 + 0.5);
  const int y_bottom = (int)((double)y1 + 0.5);
  const float r_x_1 = (float)((double)((float)b.x_left - x_1) + 0.5);
  const float r_y_1 = (float)((double)((float)b.y_top - y_1) + 0.5);
  const float r_x1 = (float)((double)(x1 - (float)x_right) + 0.5);
  const float r_y1 = (float)((double)(y1 - (float)y_bottom) + 0.5);
  b.dx = x_right - b.x_left - 1;
  b.dy = y_bottom - b.y_top - 1;
  const float fs = (float)b.scaling;
  b.A = (int)((r_x_1 * r_y_1) * fs);
  b.B = (int)((r_x1 * r_y_1) * fs);
  b.C = (int)((r_x1 * r_y1) * fs);
  b.D = (int)((r_x_1 * r_y1) * fs);
  b.wx0 = (int)(r_x_1 * fs);
  b.wy0 = (int)(r_y_1 * fs);
  b.wx1 = (int)(r_x1 * fs);
  b.wy1 = (int)(r_y1 * fs);
  return b;
}

// Rows of a box: 0 = top, 1..nMid = middle, nMid + 1 = bottom, with nMid = max(dy, 0) exactly as the reference's
// loops advance the row pointer (:450-534, integral-image path and loop path give the same weighted terms).
__device__ __forceinline__ int box_rows(const BoxParams& b) { return (b.dy > 0 ? b.dy : 0) + 2; }

// kWords: the image base and the row pitch are multiples of 4, so a row segment is read as aligned 32-bit words
// (one L1 sector lookup per 4 pixels instead of per pixel; the byte loads of scattered boxes are what bounds this
// kernel). dp4a's second operand carries one weight per byte: 1 on the interior bytes sums them, a single 1 picks the
// first / last pixel of the row - no masking, shifting or extraction instructions.
// Word view of a box: the same for every row (x_left and dx are fixed), so the weights are built once per box.
struct BoxWords {
  int w0, nw;                   // first 32-bit word of a row segment, number of words (<= kBoxWordsMax, else nw = 0: byte path)
  int a, e;                     // first and last interior byte (columns x_left+1 .. x_left+dx), counted from word 0
  unsigned selFirst, selLast;   // 0x01 on row[0] in word 0 and on row[dx+1] in word nw-1
};
constexpr int kBoxWordsMax = 5;

__device__ __forceinline__ BoxWords box_words(const BoxParams& b) {
  BoxWords k;
  const int p0 = b.x_left, pl = b.x_left + b.dx + 1;
  k.w0 = p0 >> 2;
  k.nw = (pl >> 2) - k.w0 + 1;
  k.a = (p0 & 3) + 1;
  k.e = k.a + b.dx - 1;
  k.selFirst = 1u << (8 * (p0 & 3));
  k.selLast = 1u << (8 * (pl & 3));
  if (k.nw > kBoxWordsMax) k.nw = 0;
  return k;
}

// 0x01 on the interior bytes of the first NW words of a row segment (zero past the segment): built once per box,
// and only for as many words as the row loop will load.
template <int NW>
__device__ __forceinline__ void box_weights(const BoxWords& k, unsigned (&wgt)[NW]) {
#pragma unroll
  for (int i = 0; i < NW; ++i) {
    const int lo = max(k.a - 4 * i, 0), hi = min(k.e - 4 * i, 3);
    wgt[i] = hi >= lo ? ((0x01010101u >> (8 * (3 - hi))) & (0x01010101u << (8 * lo))) : 0u;
  }
}

// Adds row r of the box to three running sums: its first pixel, its last pixel, its interior pixels.
// Word form (needs k.nw > 0). Branch-free: always five loads, at word indices clamped to the last word of the
// segment (a word past the segment has weight 0, and index 4 is always the last word). With "if (i < nw)" around
// each load the compiler emits a divergent branch per word and every load's latency is exposed in turn.
template <int NW>   // NW: loads per row, >= the longest segment of the warp's boxes
__device__ __forceinline__ void box_row_sums_words(const uint8_t* __restrict__ img, int W, const BoxParams& b,
                                                   const BoxWords& k, const unsigned (&wgt)[NW], int r,
                                                   unsigned& first, unsigned& last, unsigned& mid) {
  const unsigned* wrow = reinterpret_cast<const unsigned*>(img + (size_t)(b.y_top + r) * W) + k.w0;
  const int l = k.nw - 1;
  unsigned v[NW];
  v[0] = __ldg(wrow);
#pragma unroll
  for (int i = 1; i < NW - 1; ++i) v[i] = __ldg(wrow + min(i, l));
  if (NW > 1) v[NW - 1] = __ldg(wrow + l);   // index NW-1 is always the last word of the segment
#pragma unroll
  for (int i = 0; i < NW; ++i) mid = __dp4a(v[i], wgt[i], mid);
  first = __dp4a(v[0], k.selFirst, first);
  last = __dp4a(v[NW - 1], k.selLast, last);
}

// Byte form (unaligned image, or a segment longer than kBoxWordsMax words).
__device__ __forceinline__ void box_row_sums_bytes(const uint8_t* __restrict__ img, int W, const BoxParams& b, int r,
                                                   unsigned& first, unsigned& last, unsigned& mid) {
  const uint8_t* row = img + b.x_left + (size_t)(b.y_top + r) * W;
  for (int c = 1; c <= b.dx; ++c) mid += row[c];
  first += row[0];
  last += row[b.dx + 1];
}

template <bool kWords>
__device__ __forceinline__ void box_row_sums(const uint8_t* __restrict__ img, int W, const BoxParams& b,
                                             const BoxWords& k, int r, unsigned& first, unsigned& last, unsigned& mid) {
  if (kWords && k.nw > 0) {
    unsigned wgt[kBoxWordsMax];
    box_weights(k, wgt);
    box_row_sums_words(img, W, b, k, wgt, r, first, last, mid);
  } else {
    box_row_sums_bytes(img, W, b, r, first, last, mid);
  }
}

// The two lower corner terms. Reference quirk kept on purpose: on the integral-image path (dx + dy > 2) the two lower
// corner pixels are fetched with "ptr += dy * imagecols + 1" (:459), i.e. from row y_top + dy and one column to the
// right; the loop path uses the bottom row's own first / last pixel.
__device__ __forceinline__ int box_lower_corners(const uint8_t* __restrict__ img, int W, const BoxParams& b, int first,
                                                 int last) {
  if (b.dx + b.dy > 2) {
    const uint8_t* q = img + b.x_left + (size_t)b.y_top * W + (b.dx + 1) + (size_t)b.dy * W + 1;
    return b.C * (int)q[0] + b.D * (int)q[-(b.dx + 1)];
  }
  return b.D * first + b.C * last;
}

// One row's weighted contribution (used where the rows of a box are spread over lanes).
template <bool kWords>
__device__ __forceinline__ int box_row(const uint8_t* __restrict__ img, int W, const BoxParams& b, const BoxWords& k,
                                       int r) {
  unsigned uf = 0, ul = 0, um = 0;
  box_row_sums<kWords>(img, W, b, k, r, uf, ul, um);
  const int first = (int)uf, last = (int)ul, mid = (int)um;
  if (r == 0) return b.A * first + b.B * last + mid * b.wy0;
  if (r < box_rows(b) - 1) return first * b.wx0 + last * b.wx1 + mid * b.scaling;
  return mid * b.wy1 + box_lower_corners(img, W, b, first, last);
}

// A whole box by one lane: the middle rows share their three weights, so their pixels are summed first and weighted
// once (the reference's int32 result is a linear combination of pixel sums; the order of the additions is free).
template <bool kWords>
__device__ __forceinline__ int smoothed_intensity(const uint8_t* __restrict__ img, int W, int H, float key_x,
                                                  float key_y, PatternPoint p0, bool warped, const float* warp,
                                                  float warpScale) {
  const BoxParams b = box_params(img, W, H, key_x, key_y, p0, warped, warp, warpScale);
  if (b.mode != 2) return b.value;
  const int rows = box_rows(b);
  const BoxWords k = box_words(b);
  unsigned tf = 0, tl = 0, tm = 0, mf = 0, ml = 0, mm = 0, bf = 0, bl = 0, bm = 0;
  if (kWords && k.nw > 0) {   // decided once per box, so that the row loop is straight-line code
    // loads per row = the longest segment among the boxes the warp is summing right now (warp-uniform: no divergence)
    const int nwMax = __reduce_max_sync(__activemask(), k.nw);
    auto sum_rows = [&](auto NWc) {
      constexpr int NW = decltype(NWc)::value;
      unsigned wgt[NW];
      box_weights(k, wgt);
      box_row_sums_words(img, W, b, k, wgt, 0, tf, tl, tm);
      box_row_sums_words(img, W, b, k, wgt, rows - 1, bf, bl, bm);
#pragma unroll 2
      for (int r = 1; r < rows - 1; ++r) box_row_sums_words(img, W, b, k, wgt, r, mf, ml, mm);
    };
    if (nwMax <= 2) sum_rows(std::integral_constant<int, 2>{});
    else if (nwMax == 3) sum_rows(std::integral_constant<int, 3>{});
    else if (nwMax == 4) sum_rows(std::integral_constant<int, 4>{});
    else sum_rows(std::integral_constant<int, 5>{});
  } else {
    box_row_sums_bytes(img, W, b, 0, tf, tl, tm);
    box_row_sums_bytes(img, W, b, rows - 1, bf, bl, bm);
    for (int r = 1; r < rows - 1; ++r) box_row_sums_bytes(img, W, b, r, mf, ml, mm);
  }
  int ret = b.A * (int)tf + b.B * (int)tl + (int)tm * b.wy0;
  ret += (int)mf * b.wx0 + (int)ml * b.wx1 + (int)mm * b.scaling;
  ret += (int)bm * b.wy1 + box_lower_corners(img, W, b, (int)bf, (int)bl);
  return ret / b.scaling2;
}

// All 66 samples of one keypoint by one warp: samples 0..63 one per lane in two passes; the last two (outermost
// ring, the largest boxes) row-wise by a half-warp each instead of a third pass with 30 idle lanes.
static_assert(kPoints == 66, "sample_pattern splits 66 samples as 2 x 32 + 2");
template <bool kWords>
__device__ __forceinline__ void sample_pattern(const uint8_t* __restrict__ img, int W, int H, float x, float y,
                                               const PatternPoint* __restrict__ pat, bool warped, const float* warp,
                                               float warpScale, int lane, int* values) {
  for (int i = lane; i < 64; i += 32) values[i] = smoothed_intensity<kWords>(img, W, H, x, y, pat[i], warped, warp, warpScale);
  const int s = 64 + (lane >> 4), j = lane & 15;
  const BoxParams b = box_params(img, W, H, x, y, pat[s], warped, warp, warpScale);
  int part = 0;
  if (b.mode == 2) {
    const int rows = box_rows(b);
    const BoxWords k = box_words(b);
    for (int r = j; r < rows; r += 16) part += box_row<kWords>(img, W, b, k, r);
  }
  for (int o = 8; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (j == 0) values[s] = b.mode == 2 ? part / b.scaling2 : b.value;
}

// cv::normalize(Vec3f): norm in double, scale in double, cast to float (OpenCV matx.hpp).
__device__ __forceinline__ void cv_normalize3(const float v[3], float out[3]) {
  double s = 0.0;
  for (int i = 0; i < 3; ++i) {
    double e = (double)v[i];
    s += e * e;
  }
  const double nv = sqrt(s);
  const double inv = nv != 0.0 ? 1.0 / nv : 0.0;
  for (int i = 0; i < 3; ++i) out[i] = (float)((double)v[i] * inv);
}
__device__ __forceinline__ void cross3f(const float a[3], const float b[3], float o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ float dot3f(const float a[3], const float b[3]) {
  float s = 0.0f;
  s += a[0] * b[0];
  s += a[1] * b[1];
  s += a[2] * b[2];
  return s;
}
// Jacobi helper of cv::eigen (OpenCV lapack.cpp, JacobiImpl_<float>::hypot)
__device__ __forceinline__ float jhypot(float a, float b) {
  a = fabsf(a);
  b = fabsf(b);
  if (a > b) {
    b /= a;
    return a * sqrtf(1 + b * b);
  }
  if (b > 0) {
    a /= b;
    return b * sqrtf(1 + a * a);
  }
  return 0.0f;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32, 4)
describe_kernel(const Slot* __restrict__ slots, const PatternPoint* __restrict__ pattern,
                const uchar2* __restrict__ short_pairs, const LongPair* __restrict__ long_pairs, DescribeArgs a) {
  const Slot& slot = slots[blockIdx.z];
  const uint8_t* __restrict__ img = slot.image;
  okc_keypoint* __restrict__ kps = slot.kp_desc;
  uint8_t* __restrict__ desc = slot.desc;
  const float* __restrict__ rays = slot.rays;
  const float* __restrict__ jac = slot.jac;
  __shared__ int s_values[kWarpsPerBlock][kPoints + 2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * kWarpsPerBlock + warp;
  if (k >= slot.counters[2]) return;
  int* values = s_values[warp];
  const int W = a.width, H = a.height;
  const bool words = (W & 3) == 0 && (reinterpret_cast<uintptr_t>(img) & 3) == 0;   // uniform per slot

  okc_keypoint kp = kps[k];
  const float x = kp.x, y = kp.y;

  // ---- camera-aware warp (:654-725); every lane computes it redundantly ----
  bool warped = false, directional = false;
  float warp4[4] = {0.f, 0.f, 0.f, 0.f};
  float sigmaScale = 1.0f;
  if (slot.camera_aware) {
    const int x0 = (int)floorf(x), y0 = (int)floorf(y);
    const float dx = x - (float)x0, dy = y - (float)y0;
    const float w_tl = (1.0f - dx) * (1.0f - dy);
    const float w_tr = dx * (1.0f - dy);
    const float w_bl = (1.0f - dx) * dy;
    const float w_br = dx * dy;
    float d_tl[3], d_tr[3], d_bl[3], d_br[3];
    const float* r0 = rays + ((size_t)y0 * W + x0) * 3;
    const float* r1 = r0 + (size_t)W * 3;
    for (int i = 0; i < 3; ++i) {
      d_tl[i] = __ldg(r0 + i);
      d_tr[i] = __ldg(r0 + 3 + i);
      d_bl[i] = __ldg(r1 + i);
      d_br[i] = __ldg(r1 + 3 + i);
    }
    if ((double)dot3f(d_tl, d_tl) > 1e-12 && (double)dot3f(d_tr, d_tr) > 1e-12 && (double)dot3f(d_bl, d_bl) > 1e-12 &&
        (double)dot3f(d_br, d_br) > 1e-12) {
      float dir[3];
      for (int i = 0; i < 3; ++i) dir[i] = ((w_tl * d_tl[i] + w_tr * d_tr[i]) + w_bl * d_bl[i]) + w_br * d_br[i];
      float eu[3], ev[3], t[3];
      cross3f(slot.dir, dir, eu);
      if ((double)(eu[0] * eu[0] + eu[1] * eu[1] + eu[2] * eu[2]) > 0.01) directional = true;
      cv_normalize3(eu, t);
      for (int i = 0; i < 3; ++i) eu[i] = t[i] * slot.inv_f;
      cross3f(dir, eu, t);
      cv_normalize3(t, ev);
      for (int i = 0; i < 3; ++i) ev[i] = ev[i] * slot.inv_f;
      const int xm = (int)roundf(x), ym = (int)roundf(y);
      const float* jp = jac + ((size_t)ym * W + xm) * 6;
      float J[6];
      for (int i = 0; i < 6; ++i) J[i] = __ldg(jp + i);
      // warp = J(2x3) * [eu ev](3x2) through cv::gemm: double accumulation in k order, cast to float
      for (int r = 0; r < 2; ++r) {
        double s0 = 0.0, s1 = 0.0;
        for (int c = 0; c < 3; ++c) {
          s0 += (double)J[r * 3 + c] * (double)eu[c];
          s1 += (double)J[r * 3 + c] * (double)ev[c];
        }
        warp4[r * 2 + 0] = (float)s0;
        warp4[r * 2 + 1] = (float)s1;
      }
      // cv::eigen(2x2 float): one Jacobi rotation on the upper off-diagonal; only the eigenvalues matter
      float W0 = warp4[0], W1 = warp4[3];
      const float p = warp4[1];
      if (fabsf(p) > 1.1920928955078125e-7f) {
        const float yy = (W1 - W0) * 0.5f;
        float t2 = fabsf(yy) + jhypot(p, yy);
        t2 = (p / t2) * p;
        if (yy < 0) t2 = -t2;
        W0 -= t2;
        W1 += t2;
      }
      sigmaScale = (float)(0.5 * ((double)fabsf(W0) + (double)fabsf(W1)));
      if (directional) kp.angle = (float)(atan2((double)warp4[3], (double)warp4[1]) / 3.14159265358979323846 * 180.0);
      warped = true;
    }
  }

  // ---- orientation (:731-784) ----
  int theta;
  if (kp.angle == -1.0f) {
    if (!a.rotation_invariant) {
      theta = 0;
    } else {
      if (words) sample_pattern<true>(img, W, H, x, y, pattern, false, warp4, 1.0f, lane, values);
      else sample_pattern<false>(img, W, H, x, y, pattern, false, warp4, 1.0f, lane, values);
      __syncwarp();
      int direction0 = 0, direction1 = 0;
      for (int p = lane; p < kLongPairs; p += 32) {
        const LongPair lp = long_pairs[p];
        const int delta_t = values[lp.i] - values[lp.j];
        direction0 += delta_t * lp.wdx / 1024;
        direction1 += delta_t * lp.wdy / 1024;
      }
      for (int o = 16; o > 0; o >>= 1) {
        direction0 += __shfl_xor_sync(0xffffffffu, direction0, o);
        direction1 += __shfl_xor_sync(0xffffffffu, direction1, o);
      }
      __syncwarp();
      kp.angle = (float)(atan2((double)(float)direction1, (double)(float)direction0) / 3.14159265358979323846 * 180.0);
      theta = (int)((double)(1024.0f * kp.angle) / 360.0 + 0.5);
      if (theta < 0) theta += kNRot;
      if (theta >= kNRot) theta -= kNRot;
    }
  } else {
    if (!a.rotation_invariant) {
      theta = 0;
    } else {
      theta = (int)(1024.0 * ((double)kp.angle / 360.0) + 0.5);
      if (theta < 0) theta += kNRot;
      if (theta >= kNRot) theta -= kNRot;
    }
  }

  // ---- sampling in the final direction (:786-797) and bit tests (:543-568) ----
  const PatternPoint* pat = pattern + (size_t)(directional ? 0 : theta) * kPoints;
  if (words) sample_pattern<true>(img, W, H, x, y, pat, warped, warp4, sigmaScale, lane, values);
  else sample_pattern<false>(img, W, H, x, y, pat, warped, warp4, sigmaScale, lane, values);
  __syncwarp();
  uint32_t* out = reinterpret_cast<uint32_t*>(desc + (size_t)k * OKC_DESC_BYTES);
#pragma unroll
  for (int w = 0; w < kDescWords; ++w) {
    const uchar2 sp = short_pairs[w * 32 + lane];
    const unsigned bits = __ballot_sync(0xffffffffu, values[sp.x] > values[sp.y]);
    if (lane == 0) out[w] = bits;
  }
  if (lane == 0) kps[k].angle = kp.angle;
}

void launch_describe(const Slot* slots, int n, const PatternTables& t, const DescribeArgs& a, int max_kp,
                     cudaStream_t s) {
  border_filter_kernel<<<n, 1024, 0, s>>>(slots, a.width, a.height, t.border, max_kp);
  dim3 grid((max_kp + kWarpsPerBlock - 1) / kWarpsPerBlock, 1, n);
  describe_kernel<<<grid, kWarpsPerBlock * 32, 0, s>>>(slots, t.points, t.short_pairs, t.long_pairs, a);
}

}  // namespace okc
