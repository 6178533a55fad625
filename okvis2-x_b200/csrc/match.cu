// Hamming matching kernels.
//
// All of them replace loops around brisk::Hamming::PopcntofXORed(a, b, 3)
// (brisk/include/brisk/internal/hamming-inl.h:95-144): popcount(a XOR b) over 384 bits, and a strict '<'
// running minimum, so the FIRST candidate in iteration order wins ties.
//
// No tensor cores: the work is XOR + POPC on 12 x u32 per pair. Descriptors are read as uint4 (3 per row).
#include "common.cuh"
#include <cstdlib>

namespace okc {

// ----------------------------------------------------------------------------------------------
// A. plain brute force, queries in registers, map rows streamed through shared memory
// ----------------------------------------------------------------------------------------------
constexpr int kMapTile = 512;                   // map rows staged per block iteration by the gated matchers
constexpr int kMbThreads = 256;
constexpr int kMbQpt = 4;                       // queries per thread
constexpr int kMbQTile = kMbThreads * kMbQpt;   // 1024 queries per CTA
constexpr int kMbRows = 128;                    // map rows per shared-memory stage (6 KB)

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// Carry-save popcount of the first 5 words (160 bits): two full adders and a half adder turn 5 POPC into 3
// (POPC issues at a quarter of the LOP3 rate and is what bounds the plain loop). Exact.
__device__ __forceinline__ unsigned lop3_xor3(unsigned a, unsigned b, unsigned c) {
  unsigned d;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ unsigned lop3_maj(unsigned a, unsigned b, unsigned c) {
  unsigned d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ unsigned csa_popc5(unsigned x0, unsigned x1, unsigned x2, unsigned x3, unsigned x4) {
  const unsigned s1 = lop3_xor3(x0, x1, x2), c1 = lop3_maj(x0, x1, x2);
  const unsigned s2 = lop3_xor3(s1, x3, x4), c2 = lop3_maj(s1, x3, x4);
  const unsigned s3 = c1 ^ c2, c3 = c1 & c2;
  return __popc(s2) + 2u * __popc(s3) + 4u * __popc(c3);
}

// best[q] (u64, pre-set to ~0) receives min over rows of (dist << 32 | global row index) with dist < threshold.
// PREFIX = number of leading 32-bit words summed before the early-out test "partial >= current best"
// (exact: the full distance can only be larger). PREFIX = 12 disables the early-out.
template <int PREFIX>
__global__ void __launch_bounds__(kMbThreads)
hamming_best_kernel(const uint4* __restrict__ queries, int nq, const uint4* __restrict__ map, long long nm,
                    int rowsPerChunk, unsigned threshold, unsigned long long indexOffset,
                    unsigned long long* __restrict__ best) {
  __shared__ uint4 stage[2][kMbRows * 3];
  const int tid = threadIdx.x;
  const int qbase = blockIdx.y * kMbQTile;
  const long long row0 = (long long)blockIdx.x * rowsPerChunk;
  long long rowsLeft = nm - row0;
  if (rowsLeft > rowsPerChunk) rowsLeft = rowsPerChunk;
  if (rowsLeft <= 0) return;

  unsigned q[kMbQpt][kDescWords];
  unsigned bestd[kMbQpt], besti[kMbQpt];
#pragma unroll
  for (int j = 0; j < kMbQpt; ++j) {
    const int qi = qbase + j * kMbThreads + tid;
    bestd[j] = qi < nq ? threshold : 0u;  // 0: nothing can be strictly smaller -> lane idles
    besti[j] = 0xffffffffu;
    const int qs = qi < nq ? qi : 0;
#pragma unroll
    for (int w = 0; w < 3; ++w) {
      const uint4 v = __ldg(queries + (size_t)qs * 3 + w);
      q[j][4 * w + 0] = v.x;
      q[j][4 * w + 1] = v.y;
      q[j][4 * w + 2] = v.z;
      q[j][4 * w + 3] = v.w;
    }
  }

  const int nStages = (int)((rowsLeft + kMbRows - 1) / kMbRows);
  auto issue = [&](int s) {
    const long long r = row0 + (long long)s * kMbRows;
    long long cnt = rowsLeft - (long long)s * kMbRows;
    if (cnt > kMbRows) cnt = kMbRows;
    const uint4* src = map + r * 3;
    uint4* dst = stage[s & 1];
    for (int i = tid; i < (int)cnt * 3; i += kMbThreads) cp_async16(dst + i, src + i);
    cp_async_commit();
  };
  issue(0);
  for (int s = 0; s < nStages; ++s) {
    if (s + 1 < nStages) {
      issue(s + 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    long long cnt = rowsLeft - (long long)s * kMbRows;
    if (cnt > kMbRows) cnt = kMbRows;
    const uint4* rows = stage[s & 1];
    const unsigned rowIdx0 = (unsigned)(s * kMbRows);
#pragma unroll 2
    for (int r = 0; r < (int)cnt; ++r) {
      const uint4 m0 = rows[r * 3 + 0], m1 = rows[r * 3 + 1], m2 = rows[r * 3 + 2];
      const unsigned m[kDescWords] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
#pragma unroll
      for (int j = 0; j < kMbQpt; ++j) {
        unsigned d = 0;
        if (PREFIX == 5) {
          d = csa_popc5(q[j][0] ^ m[0], q[j][1] ^ m[1], q[j][2] ^ m[2], q[j][3] ^ m[3], q[j][4] ^ m[4]);
        } else {
#pragma unroll
          for (int w = 0; w < PREFIX; ++w) d += __popc(q[j][w] ^ m[w]);
        }
        if (d < bestd[j]) {
#pragma unroll
          for (int w = PREFIX; w < kDescWords; ++w) d += __popc(q[j][w] ^ m[w]);
          if (d < bestd[j]) {
            bestd[j] = d;
            besti[j] = rowIdx0 + r;
          }
        }
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < kMbQpt; ++j) {
    const int qi = qbase + j * kMbThreads + tid;
    if (qi < nq && besti[j] != 0xffffffffu) {
      const unsigned long long key =
          ((unsigned long long)bestd[j] << 32) | ((unsigned long long)(row0 + besti[j]) + indexOffset);
      atomicMin(best + qi, key);
    }
  }
}

// ----------------------------------------------------------------------------------------------
// A'. the default sweep kernel: 2-POPC lower bound over the first 6 words, exact distance only when it passes
// ----------------------------------------------------------------------------------------------
// The CSA-prefix kernel above keeps the ALU pipe (LOP3/ISETP, 77 % busy) and the XU pipe (POPC, 73 % busy) equally
// loaded (profiles/r01_ncu_v3), so only fewer operations per pair help. Per bit position let n = number of differing
// words among x0..x5 (x_i = q_i ^ m_i). With ones = parity(n), c1 = maj(x0,x1,x2), c2 = maj(x3,x4,x5),
// c3 = (x0^x1^x2) & (x3^x4^x5):   n >= ones + 2*(c1|c2|c3), hence
//     LB = popc(ones) + 2*popc(c1|c2|c3) <= popcount of the 6-word prefix <= full distance,
// and "LB >= running best" rejects exactly (the reference's update is a strict '<', Frontend.cpp:334-339).
// For unrelated descriptors LB ~ 73 +- 3.9 against a threshold of 60: 4e-4 of the pairs reach the exact path.
// XORs of q with m are folded into the adders through per-side precomputed combinations
// (maj(a,b,c) = ((a^b)&(a^c))^a with a^b = (q0^q1)^(m0^m1)): 10 LOP3 + 2 POPC per pair instead of 11 + 3.
// The map-side combinations are built once per staged row in shared memory.
constexpr int kLbWords = 12;  // 9 used: {m0, m0^m1, m0^m2, m0^m1^m2, m3, m3^m4, m3^m5, m3^m4^m5, m0^..^m5}, padded to 48 B

__device__ __forceinline__ unsigned lop3_and_xor(unsigned a, unsigned b, unsigned c) {  // a & (b ^ c)
  unsigned d;
  asm("lop3.b32 %0, %1, %2, %3, 0x60;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ unsigned lop3_or3(unsigned a, unsigned b, unsigned c) {
  unsigned d;
  asm("lop3.b32 %0, %1, %2, %3, 0xFE;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ void lb_combine(const unsigned w[6], unsigned c[9]) {
  c[0] = w[0];
  c[1] = w[0] ^ w[1];
  c[2] = w[0] ^ w[2];
  c[3] = c[1] ^ w[2];
  c[4] = w[3];
  c[5] = w[3] ^ w[4];
  c[6] = w[3] ^ w[5];
  c[7] = c[5] ^ w[5];
  c[8] = c[3] ^ c[7];
}

// z += LB << (8 * J) for one (query, map row) pair; q and (ma, mb, ms) hold the combinations of lb_combine.
template <int J>
__device__ __forceinline__ unsigned lb_accum(unsigned z, const unsigned (&q)[9], const uint4& ma, const uint4& mb, unsigned ms) {
  const unsigned t1 = q[1] ^ ma.y;                                   // x0 ^ x1
  const unsigned c1 = lop3_xor3(lop3_and_xor(t1, q[2], ma.z), q[0], ma.x);
  const unsigned u1 = q[5] ^ mb.y;                                   // x3 ^ x4
  const unsigned c2 = lop3_xor3(lop3_and_xor(u1, q[6], mb.z), q[4], mb.x);
  const unsigned s1 = q[3] ^ ma.w;                                   // x0 ^ x1 ^ x2
  const unsigned c3 = lop3_and_xor(s1, q[7], mb.w);
  const unsigned hi = lop3_or3(c1, c2, c3);
  const unsigned ones = q[8] ^ ms;
  asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(z) : "r"((unsigned)__popc(hi)), "n"(2u << (8 * J)));
  asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(z) : "r"((unsigned)__popc(ones)), "n"(1u << (8 * J)));
  return z;
}

__global__ void __launch_bounds__(kMbThreads, 3)
hamming_best_lb_kernel(const uint4* __restrict__ queries, int nq, const uint4* __restrict__ map, long long nm,
                       int rowsPerChunk, unsigned threshold, unsigned long long indexOffset,
                       unsigned long long* __restrict__ best) {
  __shared__ uint4 stage[2][kMbRows * 3];          // raw rows (exact path)
  __shared__ uint4 comb[kMbRows * (kLbWords / 4)]; // combinations of the first 6 words
  const int tid = threadIdx.x;
  const int qbase = blockIdx.y * kMbQTile;
  const long long row0 = (long long)blockIdx.x * rowsPerChunk;
  long long rowsLeft = nm - row0;
  if (rowsLeft > rowsPerChunk) rowsLeft = rowsPerChunk;
  if (rowsLeft <= 0) return;

  unsigned qc[kMbQpt][9];
  unsigned bestd[kMbQpt], besti[kMbQpt];
#pragma unroll
  for (int j = 0; j < kMbQpt; ++j) {
    const int qi = qbase + j * kMbThreads + tid;
    bestd[j] = qi < nq ? threshold : 0u;  // 0: nothing can be strictly smaller -> lane idles
    besti[j] = 0xffffffffu;
    const int qs = qi < nq ? qi : 0;
    const uint4 v0 = __ldg(queries + (size_t)qs * 3), v1 = __ldg(queries + (size_t)qs * 3 + 1);
    const unsigned w[6] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y};
    lb_combine(w, qc[j]);
  }
  unsigned zbase = 0;
#pragma unroll
  for (int j = 0; j < kMbQpt; ++j) zbase |= (128u - (bestd[j] < 128u ? bestd[j] : 128u)) << (8 * j);

  const int nStages = (int)((rowsLeft + kMbRows - 1) / kMbRows);
  auto issue = [&](int s) {
    const long long r = row0 + (long long)s * kMbRows;
    long long cnt = rowsLeft - (long long)s * kMbRows;
    if (cnt > kMbRows) cnt = kMbRows;
    const uint4* src = map + r * 3;
    uint4* dst = stage[s & 1];
    for (int i = tid; i < (int)cnt * 3; i += kMbThreads) cp_async16(dst + i, src + i);
    cp_async_commit();
  };
  issue(0);
  for (int s = 0; s < nStages; ++s) {
    cp_async_wait<0>();
    __syncthreads();  // stage s landed; every thread is done with comb of stage s-1
    long long cnt = rowsLeft - (long long)s * kMbRows;
    if (cnt > kMbRows) cnt = kMbRows;
    const uint4* rows = stage[s & 1];
    if (tid < (int)cnt) {
      const uint4 a = rows[tid * 3], b = rows[tid * 3 + 1];
      const unsigned w[6] = {a.x, a.y, a.z, a.w, b.x, b.y};
      unsigned c[9];
      lb_combine(w, c);
      comb[tid * 3 + 0] = make_uint4(c[0], c[1], c[2], c[3]);
      comb[tid * 3 + 1] = make_uint4(c[4], c[5], c[6], c[7]);
      comb[tid * 3 + 2] = make_uint4(c[8], 0u, 0u, 0u);
    }
    if (s + 1 < nStages) issue(s + 1);  // stage[(s+1)&1] was last read in iteration s-1, before the barrier above
    __syncthreads();
    const unsigned rowIdx0 = (unsigned)(s * kMbRows);
#pragma unroll 2
    for (int r = 0; r < (int)cnt; ++r) {
      const uint4 ma = comb[r * 3 + 0], mb = comb[r * 3 + 1];
      const unsigned ms = reinterpret_cast<const unsigned*>(comb + r * 3 + 2)[0];
      // byte j of z = 128 + LB_j - min(best_j, 128): bit 7 clear <=> LB_j < best_j (LB <= 96, so no carries between
      // bytes). The sums run on the FMA pipe (IMAD), leaving the ALU pipe to the LOP3s; one LOP3 tests all four pairs.
      unsigned z = zbase;
      z = lb_accum<0>(z, qc[0], ma, mb, ms);
      z = lb_accum<1>(z, qc[1], ma, mb, ms);
      z = lb_accum<2>(z, qc[2], ma, mb, ms);
      z = lb_accum<3>(z, qc[3], ma, mb, ms);
      if ((~z & 0x80808080u) != 0u) {
        const uint4 m0 = rows[r * 3 + 0], m1 = rows[r * 3 + 1], m2 = rows[r * 3 + 2];
#pragma unroll
        for (int j = 0; j < kMbQpt; ++j) {
          if (((z >> (8 * j + 7)) & 1u) == 0u) {
            const uint4* qp = queries + (size_t)(qbase + j * kMbThreads + tid) * 3;  // bestd > 0 implies qi < nq
            const uint4 q0 = __ldg(qp), q1 = __ldg(qp + 1), q2 = __ldg(qp + 2);
            const unsigned d = __popc(q0.x ^ m0.x) + __popc(q0.y ^ m0.y) + __popc(q0.z ^ m0.z) + __popc(q0.w ^ m0.w) +
                               __popc(q1.x ^ m1.x) + __popc(q1.y ^ m1.y) + __popc(q1.z ^ m1.z) + __popc(q1.w ^ m1.w) +
                               __popc(q2.x ^ m2.x) + __popc(q2.y ^ m2.y) + __popc(q2.z ^ m2.z) + __popc(q2.w ^ m2.w);
            if (d < bestd[j]) {
              bestd[j] = d;
              besti[j] = rowIdx0 + r;
              zbase = (zbase & ~(0xffu << (8 * j))) | ((128u - (d < 128u ? d : 128u)) << (8 * j));
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kMbQpt; ++j) {
    const int qi = qbase + j * kMbThreads + tid;
    if (qi < nq && besti[j] != 0xffffffffu) {
      const unsigned long long key =
          ((unsigned long long)bestd[j] << 32) | ((unsigned long long)(row0 + besti[j]) + indexOffset);
      atomicMin(best + qi, key);
    }
  }
}

static int pick_rows_per_chunk(int nq, long long nm, int ctasPerSm) {
  // Aim at 16 waves of 148 SMs x resident CTAs (measured: 2 waves 1189, 4 -> 1270, 16 -> 1350 Gpairs/s on the
  // 10k x 1M sweep: the tail of the last wave is what small grids lose), chunks a multiple of the stage size.
  const long long qTiles = (nq + kMbQTile - 1) / kMbQTile;
  const int waves = 16;
  long long wantChunks = (148LL * ctasPerSm * waves + qTiles - 1) / qTiles;
  long long rows = (nm + wantChunks - 1) / wantChunks;
  rows = ((rows + kMbRows - 1) / kMbRows) * kMbRows;
  if (rows < kMbRows) rows = kMbRows;
  if (rows > 65536) rows = 65536;
  return (int)rows;
}

__global__ void fill_keys_kernel(unsigned long long* __restrict__ best, int n, unsigned long long v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) best[i] = v;
}

// noMatch: the key of "no row below the threshold" (~0 = OKC_NO_MATCH, or INT64_MAX for keys that are min-reduced as
// signed integers); any value above every real key (dist <= 384 in bits 32..40) works with the kernels' atomicMin
int launch_hamming_best(const void* queries, int nq, const void* map, long long nm, unsigned threshold,
                        unsigned long long indexOffset, unsigned long long* best, cudaStream_t s,
                        unsigned long long noMatch) {
  if (nq <= 0) return 0;
  if (noMatch == ~0ull) cudaMemsetAsync(best, 0xff, sizeof(unsigned long long) * (size_t)nq, s);
  else fill_keys_kernel<<<(nq + 255) / 256, 256, 0, s>>>(best, nq, noMatch);
  if (nm <= 0) return 0;
  const bool plain = g_knobs.match_plain != 0;   // tests: the plain 12-word sweep as the A/B baseline
  const int rowsPerChunk = pick_rows_per_chunk(nq, nm, plain ? 2 : 3);
  dim3 grid((unsigned)((nm + rowsPerChunk - 1) / rowsPerChunk), (unsigned)((nq + kMbQTile - 1) / kMbQTile));
  const uint4* q4 = static_cast<const uint4*>(queries);
  const uint4* m4 = static_cast<const uint4*>(map);
  if (plain) hamming_best_kernel<12><<<grid, kMbThreads, 0, s>>>(q4, nq, m4, nm, rowsPerChunk, threshold, indexOffset, best);
  else hamming_best_lb_kernel<<<grid, kMbThreads, 0, s>>>(q4, nq, m4, nm, rowsPerChunk, threshold, indexOffset, best);
  return (int)(grid.x * grid.y);
}

// ----------------------------------------------------------------------------------------------
// shared helpers for the small gated matchers: one warp per keypoint, lanes stride over the other side
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_desc(const uint8_t* base, size_t row, unsigned d[kDescWords]) {
  const uint4* p = reinterpret_cast<const uint4*>(base) + row * 3;
  const uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  d[0] = a.x; d[1] = a.y; d[2] = a.z; d[3] = a.w;
  d[4] = b.x; d[5] = b.y; d[6] = b.z; d[7] = b.w;
  d[8] = c.x; d[9] = c.y; d[10] = c.z; d[11] = c.w;
}
__device__ __forceinline__ unsigned hamming384(const unsigned a[kDescWords], const unsigned b[kDescWords]) {
  unsigned d = 0;
#pragma unroll
  for (int w = 0; w < kDescWords; ++w) d += __popc(a[w] ^ b[w]);
  return d;
}

// ----------------------------------------------------------------------------------------------
// B. 3D-2D map matching with the pixel gate, Frontend::matchToMapByThread (Frontend.cpp:1744-1827)
// ----------------------------------------------------------------------------------------------
// One warp per keypoint k. Rows m are visited in ascending order (= landmark-major, descriptor-minor, the
// reference's iteration order for a fixed k). Every improvement of the running minimum is an "event"; the
// events (k, m) are appended to a list so that the host adapter can rebuild ctrs[] / reprErrors[] in the
// reference's accumulation order.
__device__ __forceinline__ void
match_to_map_body(const uint8_t* __restrict__ queries, const double* __restrict__ kp_xy,
                    const okc_keypoint* __restrict__ kp_f, const uint8_t* __restrict__ use, int nkMax,
                    const int* __restrict__ nkDev, const uint8_t* __restrict__ map,
                    const int* __restrict__ owner, const uint8_t* __restrict__ is3d, int nm,
                    const double* __restrict__ proj_xy, double threshSq,
                    double hammingThreshold, double* __restrict__ bestDist, int* __restrict__ bestOwner,
                    unsigned long long* __restrict__ events, int eventSlab,
                    unsigned long long* __restrict__ eventsOverflow, int eventCap, int* __restrict__ eventCount) {
  // events [0, eventSlab) go to `events` (the part of the output block that always travels to the host), the rest up
  // to eventCap to `eventsOverflow` (fetched only when needed)
  // The 8 warps (= 8 keypoints) of a block walk the same map rows: each tile of kMapTile rows has its landmark slot
  // and projection staged in shared memory once per block (one coalesced owner load + one gather per row), so the
  // per-pair work of a warp is a shared-memory read, the FP64 gate, and - for the ~1 % of pairs inside the gate - the
  // 48-byte row and 12 popcounts.
  __shared__ double s_px[kMapTile], s_py[kMapTile];
  __shared__ int s_owner[kMapTile];
  const int lane = threadIdx.x & 31;
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  // nkDev: the keypoint count is still on the device (frame pipeline: queries are the camera's own descriptors)
  const int nk = nkDev ? min(*nkDev, nkMax) : nkMax;
  if ((int)((blockIdx.x * blockDim.x) >> 5) >= nk) return;   // whole block past the last keypoint
  const bool active = k < nk && (!use || use[k]);
  double best = hammingThreshold;
  int bestO = -1;
  unsigned q[kDescWords];
  double kx = 0.0, ky = 0.0;
  if (active) {
    load_desc(queries, k, q);
    // cv::KeyPoint floats widen exactly to the Eigen::Vector2d the reference compares with (Frontend.cpp:1787-1790)
    kx = kp_f ? (double)kp_f[k].x : kp_xy[2 * k];
    ky = kp_f ? (double)kp_f[k].y : kp_xy[2 * k + 1];
  }
  for (int tile = 0; tile < nm; tile += kMapTile) {
    __syncthreads();
    for (int i = threadIdx.x; i < kMapTile; i += blockDim.x) {
      const int m = tile + i;
      if (m < nm) {
        const int o = owner[m];
        s_owner[i] = o;
        // a row of a landmark that is not 3D yet never passes the gate: matchToMapByThread skips it (Frontend.cpp:1779)
        const bool skip = is3d && !is3d[o];
        s_px[i] = skip ? INFINITY : proj_xy[2 * o];
        s_py[i] = skip ? INFINITY : proj_xy[2 * o + 1];
      }
    }
    __syncthreads();
    if (!active) continue;
    const int cnt = min(kMapTile, nm - tile);
    for (int base = 0; base < cnt; base += 32) {
      const int i = base + lane;
      double d = 1e9;
      if (i < cnt) {
        const double rx = s_px[i] - kx, ry = s_py[i] - ky;
        if (!(rx * rx + ry * ry > threshSq)) {
          unsigned r[kDescWords];
          load_desc(map, tile + i, r);
          d = (double)hamming384(q, r);
        }
      }
      unsigned pending = __ballot_sync(0xffffffffu, d < best);
      while (pending) {  // rare: replay the candidates of this group in row order
        const int src = __ffs(pending) - 1;
        pending &= pending - 1;
        const double dc = __shfl_sync(0xffffffffu, d, src);
        if (dc < best) {
          best = dc;
          bestO = s_owner[base + src];
          if (lane == 0) {
            const int e = atomicAdd(eventCount, 1);
            const unsigned long long ev = ((unsigned long long)(unsigned)(tile + base + src) << 32) | (unsigned)k;
            if (e < eventSlab) events[e] = ev;
            else if (e < eventCap) eventsOverflow[e - eventSlab] = ev;
          }
        }
      }
    }
  }
  if (k >= nk) return;
  if (lane == 0) {
    bestDist[k] = best;
    bestOwner[k] = bestO;
  }
}

__global__ void __launch_bounds__(256)
match_to_map_kernel(const uint8_t* __restrict__ queries, const double* __restrict__ kp_xy,
                    const okc_keypoint* __restrict__ kp_f, const uint8_t* __restrict__ use, int nkMax,
                    const int* __restrict__ nkDev, const uint8_t* __restrict__ map, const int* __restrict__ owner,
                    int nm, const double* __restrict__ proj_xy, double threshSq, double hammingThreshold,
                    double* __restrict__ bestDist, int* __restrict__ bestOwner, unsigned long long* __restrict__ events,
                    int eventCap, int* __restrict__ eventCount) {
  match_to_map_body(queries, kp_xy, kp_f, use, nkMax, nkDev, map, owner, nullptr, nm, proj_xy, threshSq,
                    hammingThreshold, bestDist, bestOwner, events, eventCap, nullptr, eventCap, eventCount);
}

// Frame-pipeline form: one launch for all slots of a batch; queries are the slot's own descriptors / keypoints of
// the frame just described (count still on the device), every keypoint unassociated.
__global__ void __launch_bounds__(256)
match_to_map_slots_kernel(const Slot* __restrict__ slots, int maxKp, double hammingThreshold) {
  const Slot& slot = slots[blockIdx.z];
  if (slot.map == nullptr) return;  // no map match requested for this slot
  match_to_map_body(slot.desc, nullptr, slot.kp_desc, nullptr, maxKp, slot.counters + 2, slot.map, slot.owner, slot.is3d,
                    slot.nm_dev ? min(*slot.nm_dev, slot.nm) : slot.nm, slot.proj, slot.thresh_sq, hammingThreshold,
                    slot.best_dist, slot.best_owner, slot.events, slot.event_cap, nullptr, slot.event_cap,
                    slot.counters + 4);
}

// ----------------------------------------------------------------------------------------------
// B'. the same 3D-2D match through spatial bins of the projections (frame pipeline default)
// ----------------------------------------------------------------------------------------------
// The pixel gate of Frontend.cpp:1793-1796 keeps ~1 % of the (keypoint, row) pairs, yet the scan above evaluates it
// for all of them. bin_projections_kernel groups the rows of a slot's pool by the cell (side >= gate radius + 2 px)
// their landmark projection falls into, ascending row index inside a cell; a keypoint then only visits the 3x3 cells
// around its own. Exactness: the bins only preselect - the FP64 gate itself is evaluated unchanged on every visited
// row - and the reference's iteration order matters only for rows that can improve the running best, i.e. rows with
// dist < threshold: those few "hits" are collected and replayed in ascending row order, which reproduces best,
// owner and the improvement events (ctrs / reprErrors) of the full scan.
constexpr int kBinThreads = 1024;

__device__ __forceinline__ int bin_cell_of(double p, double pad, double cell, int n) {
  const double c = floor((p + pad) / cell);
  if (!(c >= 0.0 && c < (double)n)) return c != c || fabs(p) == INFINITY ? -2 : -1;   // -2: not finite
  return (int)c;
}

__global__ void __launch_bounds__(kBinThreads, 1) bin_projections_kernel(const Slot* __restrict__ slots) {
  const Slot& slot = slots[blockIdx.x];
  if (slot.map == nullptr || slot.bins_x == 0) return;
  __shared__ int s_cnt[kMaxBinCells];
  __shared__ int s_warp[32];
  __shared__ int s_flag;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ncells = slot.bins_x * slot.bins_y, nm = slot.nm_dev ? min(*slot.nm_dev, slot.nm) : slot.nm;
  for (int c = tid; c < kMaxBinCells; c += kBinThreads) s_cnt[c] = 0;
  if (tid == 0) s_flag = 0;
  __syncthreads();
  auto cell_of_row = [&](int m, double* px, double* py, int* o) {
    *o = slot.owner[m];
    if (slot.is3d && !slot.is3d[*o]) return -1;   // matchToMapByThread skips landmarks that are not 3D (Frontend.cpp:1779)
    *px = slot.proj[2 * *o];
    *py = slot.proj[2 * *o + 1];
    const int cx = bin_cell_of(*px, slot.bin_pad, slot.bin_cell, slot.bins_x);
    const int cy = bin_cell_of(*py, slot.bin_pad, slot.bin_cell, slot.bins_y);
    if (cx == -2 || cy == -2) return -2;
    if (cx < 0 || cy < 0) return -1;
    return cy * slot.bins_x + cx;
  };
  for (int m = tid; m < nm; m += kBinThreads) {
    double px, py;
    int o;
    const int c = cell_of_row(m, &px, &py, &o);
    if (c >= 0) atomicAdd(&s_cnt[c], 1);
    else if (c == -2) s_flag = 1;
  }
  __syncthreads();
  // exclusive scan of the cell counts: 4 cells per thread, warp scan, scan of the warp totals
  int v[4], sum = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[j] = s_cnt[tid * 4 + j];
    sum += v[j];
  }
  int incl = sum;
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int w = s_warp[lane];
    int wi = w;
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += t;
    }
    s_warp[lane] = wi - w;
  }
  __syncthreads();
  int run = s_warp[warp] + incl - sum;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = tid * 4 + j;
    if (c <= ncells) slot.bin_start[c] = run;   // entry [ncells] = number of binned rows
    s_cnt[c] = run;                             // becomes the fill cursor
    run += v[j];
  }
  if (tid == 0) slot.bin_start[ncells + 1] = s_flag;
  __syncthreads();
  for (int m = tid; m < nm; m += kBinThreads) {
    double px, py;
    int o;
    const int c = cell_of_row(m, &px, &py, &o);
    if (c >= 0) slot.bin_rows[atomicAdd(&s_cnt[c], 1)] = m;
  }
  __syncthreads();
  // ascending row index inside each cell (the fill order of the atomics is arbitrary): the lists are short
  for (int c = tid; c < ncells; c += kBinThreads) {
    const int b = slot.bin_start[c], e = s_cnt[c];
    for (int i = b + 1; i < e; ++i) {
      const int x = slot.bin_rows[i];
      int j = i - 1;
      while (j >= b && slot.bin_rows[j] > x) {
        slot.bin_rows[j + 1] = slot.bin_rows[j];
        --j;
      }
      slot.bin_rows[j + 1] = x;
    }
  }
  __syncthreads();
  const int total = slot.bin_start[ncells];
  for (int i = tid; i < total; i += kBinThreads) {
    const int m = slot.bin_rows[i];
    const int o = slot.owner[m];
    slot.bin_owner[i] = o;
    slot.bin_px[i] = slot.proj[2 * o];
    slot.bin_py[i] = slot.proj[2 * o + 1];
  }
}

void launch_bin_projections(const Slot* slots, int n, cudaStream_t s) {
  if (n > 0) bin_projections_kernel<<<n, kBinThreads, 0, s>>>(slots);
}

constexpr int kBinnedWarps = 4;
constexpr int kMaxHits = 64;   // rows inside the gate with dist < threshold, per keypoint (2 per lane); more: full scan

__global__ void __launch_bounds__(kBinnedWarps * 32, 48 / kBinnedWarps)
match_to_map_binned_kernel(const Slot* __restrict__ slots, int maxKp, double hammingThreshold) {
  const Slot& slot = slots[blockIdx.z];
  if (slot.map == nullptr) return;
  const int lane = threadIdx.x & 31;
  const int k = blockIdx.x * kBinnedWarps + (threadIdx.x >> 5);
  // The kernel is a chain of dependent loads (slot table -> counters -> keypoint -> cell bounds -> rows ->
  // descriptors); everything that only depends on the slot table is requested before the first branch. The
  // keypoint and its descriptor are read at a clamped index: the arrays hold maxKp rows, the count is not known yet.
  const int kSafe = k < maxKp ? k : maxKp - 1;
  const int ncells = slot.bins_x * slot.bins_y;
  const int nkRaw = slot.counters[2];
  const int notFinite = slot.bins_x > 0 ? slot.bin_start[ncells + 1] : 1;
  const okc_keypoint kpt = slot.kp_desc[kSafe];
  unsigned q[kDescWords];
  load_desc(slot.desc, kSafe, q);
  if (k >= min(nkRaw, maxKp)) return;
  // cv::KeyPoint floats widen exactly to the Eigen::Vector2d the reference compares with (Frontend.cpp:1787-1790)
  const double kx = (double)kpt.x, ky = (double)kpt.y;
  const double threshSq = slot.thresh_sq;

  // hits of this keypoint, two per lane: row index, distance (an integer in [0, 384]), landmark slot
  int hm[2] = {0x7fffffff, 0x7fffffff}, ho[2] = {-1, -1};
  unsigned hd[2] = {0, 0};
  int nh = 0;            // hits held by this lane
  bool overflow = false;
  const bool binned = notFinite == 0;
  if (binned) {
    const int cx = bin_cell_of(kx, slot.bin_pad, slot.bin_cell, slot.bins_x);
    const int cy = bin_cell_of(ky, slot.bin_pad, slot.bin_cell, slot.bins_y);
    if (cx >= 0 && cy >= 0) {   // (always, for a keypoint inside the image)
      // the 3 x 3 cells are three runs of consecutive cells = three ranges of the binned rows; lanes 0..5 fetch the
      // six bounds at once and the ranges are walked as one list
      const int x0 = cx > 0 ? cx - 1 : 0, x1 = cx + 1 < slot.bins_x ? cx + 1 : slot.bins_x - 1;
      int bound = 0;
      {
        const int yy = cy - 1 + (lane >> 1);
        if (lane < 6 && yy >= 0 && yy < slot.bins_y) bound = slot.bin_start[yy * slot.bins_x + ((lane & 1) ? x1 + 1 : x0)];
      }
      const int b0 = __shfl_sync(0xffffffffu, bound, 0), n0 = __shfl_sync(0xffffffffu, bound, 1) - b0;
      const int b1 = __shfl_sync(0xffffffffu, bound, 2), n1 = __shfl_sync(0xffffffffu, bound, 3) - b1;
      const int b2 = __shfl_sync(0xffffffffu, bound, 4), n2 = __shfl_sync(0xffffffffu, bound, 5) - b2;
      const int total = n0 + n1 + n2;
      for (int base = 0; base < total; base += 32) {
        const int t = base + lane;
        if (t < total) {
          const int i = t < n0 ? b0 + t : (t < n0 + n1 ? b1 + (t - n0) : b2 + (t - n0 - n1));
          const double rx = slot.bin_px[i] - kx, ry = slot.bin_py[i] - ky;
          if (!(rx * rx + ry * ry > threshSq)) {
            const int m = slot.bin_rows[i];
            unsigned r[kDescWords];
            load_desc(slot.map, m, r);
            const unsigned d = hamming384(q, r);
            if ((double)d < hammingThreshold) {
              if (nh < 2) {
                hm[nh] = m;
                hd[nh] = d;
                ho[nh] = slot.bin_owner[i];
                ++nh;
              } else {
                overflow = true;
              }
            }
          }
        }
      }
    }
  }
  double best = hammingThreshold;
  int bestO = -1;
  auto improve = [&](unsigned d, int m, int o) {   // uniform across the warp
    if ((double)d < best) {
      best = (double)d;
      bestO = o;
      if (lane == 0) {
        const int e = atomicAdd(slot.counters + 4, 1);
        if (e < slot.event_cap) slot.events[e] = ((unsigned long long)(unsigned)m << 32) | (unsigned)k;
      }
    }
  };
  if (binned && !__any_sync(0xffffffffu, overflow)) {
    // replay the hits in ascending row order (row indices are unique: a row lies in exactly one cell)
    for (;;) {
      const int mine = hm[0] < hm[1] ? hm[0] : hm[1];
      const int mmin = __reduce_min_sync(0xffffffffu, mine);
      if (mmin == 0x7fffffff) break;
      const int src = __ffs(__ballot_sync(0xffffffffu, mine == mmin)) - 1;
      const bool first = hm[0] == mmin;
      const unsigned d = __shfl_sync(0xffffffffu, first ? hd[0] : hd[1], src);
      const int o = __shfl_sync(0xffffffffu, first ? ho[0] : ho[1], src);
      if (lane == src) {
        if (first) hm[0] = 0x7fffffff;
        else hm[1] = 0x7fffffff;
      }
      improve(d, mmin, o);
    }
  } else {
    // full scan in row order (no bins for this slot, a non-finite projection, or too many hits): rare
    const int nmAll = slot.nm_dev ? min(*slot.nm_dev, slot.nm) : slot.nm;
    for (int base = 0; base < nmAll; base += 32) {
      const int m = base + lane;
      unsigned d = 0xffffffffu;
      int o = -1;
      if (m < nmAll) {
        o = slot.owner[m];
        const double rx = slot.proj[2 * o] - kx, ry = slot.proj[2 * o + 1] - ky;
        if (!(slot.is3d && !slot.is3d[o]) && !(rx * rx + ry * ry > threshSq)) {
          unsigned r[kDescWords];
          load_desc(slot.map, m, r);
          d = hamming384(q, r);
        }
      }
      unsigned pending = __ballot_sync(0xffffffffu, d != 0xffffffffu && (double)d < best);
      while (pending) {
        const int src = __ffs(pending) - 1;
        pending &= pending - 1;
        improve(__shfl_sync(0xffffffffu, d, src), base + src, __shfl_sync(0xffffffffu, o, src));
      }
    }
  }
  if (lane == 0) {
    slot.best_dist[k] = best;
    slot.best_owner[k] = bestO;
  }
}

void launch_match_to_map_binned(const Slot* slots, int n, int max_kp, double hammingThreshold, cudaStream_t s) {
  if (n <= 0 || max_kp <= 0) return;
  dim3 grid((max_kp + kBinnedWarps - 1) / kBinnedWarps, 1, n);
  match_to_map_binned_kernel<<<grid, kBinnedWarps * 32, 0, s>>>(slots, max_kp, hammingThreshold);
}

void launch_match_to_map_slots(const Slot* slots, int n, int max_kp, double hammingThreshold, cudaStream_t s) {
  if (n <= 0 || max_kp <= 0) return;
  const int warpsPerBlock = 8;
  dim3 grid((max_kp + warpsPerBlock - 1) / warpsPerBlock, 1, n);
  match_to_map_slots_kernel<<<grid, warpsPerBlock * 32, 0, s>>>(slots, max_kp, hammingThreshold);
}

void launch_match_to_map(const uint8_t* queries, const double* kp_xy, const okc_keypoint* kp_f, const uint8_t* use,
                         int nkMax, const int* nkDev, const uint8_t* map, const int* owner, int nm,
                         const double* proj_xy, double threshSq, double hammingThreshold, double* bestDist,
                         int* bestOwner, unsigned long long* events, int eventCap, int* eventCount, cudaStream_t s) {
  if (nkMax <= 0) return;
  const int warpsPerBlock = 8;
  const int blocks = (nkMax + warpsPerBlock - 1) / warpsPerBlock;
  match_to_map_kernel<<<blocks, warpsPerBlock * 32, 0, s>>>(queries, kp_xy, kp_f, use, nkMax, nkDev, map, owner, nm,
                                                           proj_xy, threshSq, hammingThreshold, bestDist, bestOwner,
                                                           events, eventCap, eventCount);
}

// ----------------------------------------------------------------------------------------------
// ctrs[] / reprErrors[] of matchToMapByThread, on the device
// ----------------------------------------------------------------------------------------------
// The reference counts every improvement of a keypoint's running best and adds its reprojection distance to a
// double accumulator, per thread slice, in ITS iteration order: landmark-major, then keypoint, then descriptor row
// (Frontend.cpp:1785-1826). The match kernel logs the improvements as (row, k) events in arbitrary order; this kernel
// sorts them by (slice, landmark, k, event index) - events of one k are logged in row order by one warp, so the
// event index orders them - and lets one thread per slice add the distances sequentially, which reproduces the
// reference's sums bit for bit. The final division (NaN for an empty slice) is done by the host at finish.
constexpr int kStatsThreads = 1024;
constexpr int kStatsSmemEvents = 8192;

__global__ void __launch_bounds__(kStatsThreads, 1) match_stats_kernel(const Slot* __restrict__ slots) {
  extern __shared__ __align__(16) unsigned long long s_keys[];  // kStatsSmemEvents keys + as many distances (128 KB)
  __shared__ int s_begin[kMaxSlices + 1];
  const Slot& slot = slots[blockIdx.x];
  if (slot.map == nullptr) return;
  const int tid = threadIdx.x;
  const int T = slot.num_slices;
  int E = slot.counters[4];
  if (E > slot.event_cap) E = slot.event_cap;  // overflow is reported by the host from counters[4]
  const int nk = slot.counters[2];
  const int segment = nk / T;
  if (tid <= T) s_begin[tid] = E;
  int npad = 1;
  while (npad < E) npad <<= 1;
  unsigned long long* keys = npad <= kStatsSmemEvents ? s_keys : slot.sort_scratch;
  // key = slice(6) | landmark(24) | k(16) | event index(18)
  for (int i = tid; i < npad; i += kStatsThreads) {
    unsigned long long key = ~0ull;
    if (i < E) {
      const unsigned long long ev = slot.events[i];
      const unsigned m = (unsigned)(ev >> 32), k = (unsigned)ev;
      unsigned t = segment == 0 ? (unsigned)(T - 1) : k / (unsigned)segment;
      if (t > (unsigned)(T - 1)) t = (unsigned)(T - 1);
      key = ((unsigned long long)t << 58) | ((unsigned long long)(unsigned)slot.owner[m] << 34) |
            ((unsigned long long)k << 18) | (unsigned long long)i;
    }
    keys[i] = key;
  }
  __syncthreads();
  for (int kk = 2; kk <= npad; kk <<= 1) {
    for (int j = kk >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < npad; i += kStatsThreads) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], b = keys[ixj];
          if ((a > b) == ((i & kk) == 0)) {
            keys[i] = b;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  // slice boundaries: first index whose slice field is >= t
  for (int i = tid; i < E; i += kStatsThreads) {
    const int t = (int)(keys[i] >> 58);
    const int tprev = i == 0 ? -1 : (int)(keys[i - 1] >> 58);
    for (int u = tprev + 1; u <= t; ++u) s_begin[u] = i;
  }
  __syncthreads();
  // reprojection distances of all events in parallel (the loads are L2 round trips: serialised inside the per-slice
  // sums they were most of this kernel's time), then the sequential sums in the reference's order from shared memory
  double* s_dist = reinterpret_cast<double*>(s_keys + kStatsSmemEvents);
  const bool distInSmem = E <= kStatsSmemEvents;
  auto event_dist = [&](unsigned long long key) {
    const unsigned o = (unsigned)(key >> 34) & 0xffffffu, k = (unsigned)(key >> 18) & 0xffffu;
    const double rx = slot.proj[2 * o] - (double)slot.kp_desc[k].x, ry = slot.proj[2 * o + 1] - (double)slot.kp_desc[k].y;
    return sqrt(rx * rx + ry * ry);
  };
  if (distInSmem)
    for (int i = tid; i < E; i += kStatsThreads) s_dist[i] = event_dist(keys[i]);
  __syncthreads();
  if (tid < T) {
    // s_begin[u] for slices after the last populated one stayed E; empty slices in between were filled above
    const int b = s_begin[tid], e = s_begin[tid + 1];
    double sum = 0.0;
    for (int i = b; i < e; ++i) sum += distInSmem ? s_dist[i] : event_dist(keys[i]);
    slot.stats[tid].count = (unsigned long long)(e - b);
    slot.stats[tid].sum = sum;
  }
}

void launch_match_stats(const Slot* slots, int n, cudaStream_t s) {
  if (n <= 0) return;
  static bool configured[64] = {};   // the attribute is per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    cudaFuncSetAttribute(match_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStatsSmemEvents * 16);
    configured[dev & 63] = true;
  }
  match_stats_kernel<<<n, kStatsThreads, kStatsSmemEvents * 16, s>>>(slots);   // keys + distances
}

// ----------------------------------------------------------------------------------------------
// FP64 geometry shared by the 2D-2D and uninitialised-landmark matchers.
// Eigen expressions are written out in plain doubles: dot of 3-vectors = (x0*y0 + x1*y1) + x2*y2,
// normalized() = v / sqrt(v.v) (element-wise division), Matrix2d::computeInverseWithCheck = |det| > 1e-12.
// ----------------------------------------------------------------------------------------------
struct V3 { double x, y, z; };
__device__ __forceinline__ V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ double norm(V3 a) { return sqrt(dot(a, a)); }
__device__ __forceinline__ V3 normalized(V3 a) {
  const double z = dot(a, a);
  if (z > 0.0) {
    const double n = sqrt(z);
    return v3(a.x / n, a.y / n, a.z / n);
  }
  return a;
}
__device__ __forceinline__ V3 matvec(const double* C, V3 v) {  // row-major 3x3
  return v3((C[0] * v.x + C[1] * v.y) + C[2] * v.z, (C[3] * v.x + C[4] * v.y) + C[5] * v.z,
            (C[6] * v.x + C[7] * v.y) + C[8] * v.z);
}
__device__ __forceinline__ V3 matTvec(const double* C, V3 v) {  // C^T * v
  return v3((C[0] * v.x + C[3] * v.y) + C[6] * v.z, (C[1] * v.x + C[4] * v.y) + C[7] * v.z,
            (C[2] * v.x + C[5] * v.y) + C[8] * v.z);
}

// triangulation::triangulateFast (okvis_frontend/src/stereo_triangulation.cpp:30-111). Returns the
// (non-normalised) homogeneous point with w = 1.
// c26 = cos(2.6 sigma), c60 = cos(6 sigma): passed in so that a caller with a per-call sigma can use the host libm's
// values (what the reference's CPU code evaluates); the per-pair callers evaluate them on the device.
__device__ V3 triangulate_fast(V3 p1, V3 e1, V3 p2, V3 e2, double c26, double c60, bool& isValid, bool& isParallel) {
  isParallel = false;
  isValid = true;
  const V3 t12 = p2 - p1;
  const double b0 = dot(t12, e1), b1 = dot(t12, e2);
  const double A00 = dot(e1, e1), A10 = dot(e1, e2), A01 = -A10, A11 = -dot(e2, e2);
  const double det = A00 * A11 - A01 * A10;
  const bool invertible = fabs(det) > 1.0e-12;
  double l0 = 0.0, l1 = 0.0;
  if (invertible) {
    const double invdet = 1.0 / det;
    const double i00 = A11 * invdet, i01 = -A01 * invdet, i10 = -A10 * invdet, i11 = A00 * invdet;
    l0 = i00 * b0 + i01 * b1;
    l1 = i10 * b0 + i11 * b1;
  }
  if (!invertible || l0 < 0.01 || l1 < 0.01) {
    isParallel = true;
    const V3 m = p1 + 0.5 * t12;
    const double tn = norm(t12);
    const V3 midpoint = m + (40.0 * (0.01 > tn ? 0.01 : tn)) * (e1 + e2);
    if (dot(e1, normalized(midpoint - p1)) < c26) isValid = false;
    if (dot(e2, normalized(midpoint - p2)) < c26) isValid = false;
    return midpoint;
  }
  const V3 xm = l0 * e1 + p1;
  const V3 xn = l1 * e2 + p2;
  const V3 s = xm + xn;
  const V3 midpoint = v3(s.x / 2.0, s.y / 2.0, s.z / 2.0);
  if (dot(e1, normalized(midpoint - p1)) < c26) isValid = false;
  if (dot(e2, normalized(midpoint - p2)) < c26) isValid = false;
  if (dot(normalized(midpoint - p2), normalized(midpoint - p1)) > c60) isParallel = true;
  return midpoint;
}

// ----------------------------------------------------------------------------------------------
// C. 2D-2D stereo matching, Frontend::matchStereo inner loops (Frontend.cpp:2259-2316)
// ----------------------------------------------------------------------------------------------
// MOTION = false: Frontend::matchStereo (Frontend.cpp:2259-2316).
// MOTION = true : the worker lambda of Frontend::matchMotionStereo (Frontend.cpp:2053-2137): k0 must be eligible
// (use0: not an initialised landmark, not observed, back-projection valid - estimator state, decided by the
// caller), sigma from size0 only, an extra e0.e1 >= 0.5 pre-gate, 0.2 m depth gates, and the triangulation quality
// acos(...) of the accepted match as an additional output.
// One launch serves every job of a batch (blockIdx.y = job): camera pairs of a multiframe, or (keyframe view, current
// image) pairs. A side is either host-uploaded arrays or a batch slot whose keypoint count is still on the device.
__device__ __forceinline__ double side_size(const StereoSide& v, int k) {
  return v.size_d ? v.size_d[k] : (double)v.kp[k].size;
}

// Work layout (shared with the uninitialised-landmark matcher below): two kernels.
//   1. filter  - every (query, row) pair's Hamming distance, the whole GPU at once: a CTA compacts the eligible ones of
//      32 consecutive queries, a warp keeps Q of them in registers and sweeps ITS chunk of the rows (grid.z), one row
//      per lane and tile, the next tile requested before this one is compared (first the 6-word prefix of every pair,
//      the exact distance only for the queued few, see filter_sweep). A pair below the threshold is appended
//      to the query's candidate list ((row << 9) | dist, kCandPerQuery entries; more -> the count says "overflow").
//      Unrelated descriptors are ~192 bits apart, so the lists hold the true match and a few look-alikes.
//   2. resolve - one thread per query runs the reference's loop body over its candidates in ascending row order (= the
//      order in which the reference visits them). A query whose list overflowed walks all rows itself with the same
//      code - results are identical either way.
// The reference's loop is sequential per query only through the running best; everything else is per pair.
constexpr int kFilterQ = 4;         // queries a filter warp keeps in registers
constexpr int kFilterRows = 512;    // rows per filter CTA (grid.z chunks)
constexpr int kCandRowShift = 9;    // candidate = (row << 9) | dist, dist <= 384
constexpr int kCandMaxRows = 1 << (32 - kCandRowShift);

// rows [lo, hi) of `rowDesc` (through `rowOf`: row index and liveness of sweep position i) against the Q queries of a
// warp; q[j] valid for j < nq, out[j] = the query's index into qDesc / cand / count.
// Two passes, because the popcount pipe (16 lanes per clock and SM) is what bounds the sweep: every pair gets the
// distance over the first 6 of its 12 words (a lower bound of the distance); the few pairs whose bound is below the
// threshold are queued in shared memory (one queue per warp) and get the exact distance later, one pair per lane.
// (Two rows per lane and step - four loads in flight - was measured: slower, 116 registers.)
constexpr int kFilterQueue = 256;   // entries per warp: a tile adds at most 32 * kFilterQ = 128
template <class RowOf>
__device__ __forceinline__ void filter_sweep(const uint8_t* __restrict__ qDesc, const uint8_t* __restrict__ rowDesc,
                                             RowOf rowOf, int lo, int hi, const unsigned (&q)[kFilterQ][kDescWords],
                                             int nq, const int (&out)[kFilterQ], int thr, unsigned* __restrict__ cand,
                                             int* __restrict__ count, uint2* __restrict__ queue) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  int queued = 0;
  auto drain = [&]() {
    __syncwarp();
    for (int e0 = 0; e0 < queued; e0 += 32) {
      const int e = e0 + lane;
      if (e < queued) {
        const uint2 it = queue[e];
        unsigned a[kDescWords], b[kDescWords];
        load_desc(qDesc, it.x, a);
        load_desc(rowDesc, it.y, b);
        const int d = (int)hamming384(a, b);
        if (d < thr) {
          const int pos = atomicAdd(count + it.x, 1);
          if (pos < kCandPerQuery) cand[(size_t)it.x * kCandPerQuery + pos] = (it.y << kCandRowShift) | (unsigned)d;
        }
      }
    }
    __syncwarp();
    queued = 0;
  };
  unsigned r[6], rNext[6];
  int m = 0, mNext = 0;
  auto fetch6 = [&](int row, unsigned (&w)[6]) {
    const uint4* p = reinterpret_cast<const uint4*>(rowDesc) + (size_t)row * 3;
    const uint4 a = __ldg(p);
    const uint2 b = __ldg(reinterpret_cast<const uint2*>(p + 1));
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y;
  };
  bool live = rowOf(lo + lane, hi, m), liveNext = false;
  if (live) fetch6(m, r);
  for (int base = lo; base < hi; base += 32) {
    liveNext = rowOf(base + 32 + lane, hi, mNext);
    if (liveNext) fetch6(mNext, rNext);
#pragma unroll
    for (int j = 0; j < kFilterQ; ++j) {
      int d = 0;
#pragma unroll
      for (int w = 0; w < 6; ++w) d += __popc(q[j][w] ^ r[w]);
      const bool pass = live && j < nq && d < thr;
      const unsigned mask = __ballot_sync(0xffffffffu, pass);
      if (mask) {
        if (pass) queue[queued + __popc(mask & lt)] = make_uint2((unsigned)out[j], (unsigned)m);
        queued += __popc(mask);
      }
    }
    if (queued > kFilterQueue - 32 * kFilterQ) drain();
#pragma unroll
    for (int w = 0; w < 6; ++w) r[w] = rNext[w];
    m = mNext;
    live = liveNext;
  }
  drain();
}

__device__ __forceinline__ bool stereo_takes_part(const StereoJob& job, int k0, bool motion) {
  // motion stereo: k0 must be eligible (Frontend.cpp:2056-2064, 2082-2084); plain stereo accepts no candidate for a k0
  // without a valid back-projection (:2275-2277), so the filter can leave it out
  return motion ? ((!job.use0 || job.use0[k0]) && job.v0.bp_valid[k0]) : job.v0.bp_valid[k0] != 0;
}
struct StereoRowOf {
  const int* list1;
  const int* oa;
  const int* ob;
  __device__ __forceinline__ bool operator()(int i, int hi, int& k1) const {
    if (i >= hi) return false;
    k1 = list1 ? list1[i] : i;
    return !(oa && oa[k1] >= 0) && !(ob && ob[k1] >= 0);
  }
};
__device__ __forceinline__ int stereo_rows(const StereoJob& job) {
  // candidates k1: every keypoint of v1, or a precompacted ascending list (the slot's keypoints without a landmark)
  const StereoSide& v1 = job.v1;
  return job.v1_list ? min(*job.v1_list_n, v1.n) : (v1.n_dev ? min(*v1.n_dev, v1.n) : v1.n);
}

template <bool MOTION>
__global__ void __launch_bounds__(128) stereo_filter_kernel(const StereoJob* __restrict__ jobs, int thr) {
  __shared__ int s_list[32];
  __shared__ int s_n;
  __shared__ uint2 s_queue[4][kFilterQueue];
  const StereoJob& job = jobs[blockIdx.y];
  const StereoSide& v0 = job.v0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n0 = v0.n_dev ? min(*v0.n_dev, v0.n) : v0.n;
  const int n1 = stereo_rows(job);
  const StereoRowOf rowOf{job.v1_list, job.v1_owner_a, job.v1_owner_b};
  // work items of the job: (chunk of 32 queries) x (chunk of kFilterRows rows), dealt round-robin to the job's CTAs
  const int kChunks = (n0 + 31) / 32, items = kChunks * ((n1 + kFilterRows - 1) / kFilterRows);
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int kBase = (item % kChunks) * 32, lo = (item / kChunks) * kFilterRows;
    const int hi = min(n1, lo + kFilterRows);
    __syncthreads();
    if (warp == 0) {
      const int k0 = kBase + lane;
      const bool use = k0 < n0 && stereo_takes_part(job, k0, MOTION);
      const unsigned bal = __ballot_sync(0xffffffffu, use);
      if (use) s_list[__popc(bal & ((1u << lane) - 1u))] = k0;
      if (lane == 0) s_n = __popc(bal);
    }
    __syncthreads();
    const int nUse = s_n;
    for (int g0 = warp * kFilterQ; g0 < nUse; g0 += 4 * kFilterQ) {
      const int nq = min(kFilterQ, nUse - g0);
      unsigned q[kFilterQ][kDescWords];
      int out[kFilterQ];
#pragma unroll
      for (int j = 0; j < kFilterQ; ++j) {
        out[j] = s_list[g0 + min(j, nq - 1)];
        load_desc(v0.desc, out[j], q[j]);
      }
      filter_sweep(v0.desc, job.v1.desc, rowOf, lo, hi, q, nq, out, thr, job.cand, job.cand_count, s_queue[warp]);
    }
  }
}

// CTAs per job of a filter launch: enough to fill the GPU a few times over, never more than the job has work items
// (the counts are only known on the device: a job's CTAs share its items round-robin, none is launched just to exit)
static int filter_ctas_per_job(int maxQueries, int maxRows, int nJobs) {
  static int sms[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (sms[dev & 63] == 0) {
    int v = 148;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    sms[dev & 63] = v;
  }
  const long long items = (long long)((maxQueries + 31) / 32) * ((maxRows + kFilterRows - 1) / kFilterRows);
  long long want = ((long long)sms[dev & 63] * 12 + nJobs - 1) / nJobs;
  if (want > items) want = items;
  return (int)(want < 1 ? 1 : want);
}

// next candidate of a query in ascending row order: the smallest packed value above `last` (lists are short: a scan)
__device__ __forceinline__ unsigned next_candidate(const unsigned* __restrict__ list, int cnt, unsigned last, bool first) {
  unsigned best = ~0u;
  for (int i = 0; i < cnt; ++i) {
    const unsigned v = list[i];
    if ((first || v > last) && v < best) best = v;
  }
  return best;
}

// resolve: one THREAD per k0. The gates are ~3 000 FP64 instructions (IEEE divisions, square roots, cos, acos) per
// candidate that beats the running best and a query has about one candidate: a warp per query would run them for one
// lane's benefit.
template <bool MOTION>
__global__ void __launch_bounds__(128)
match_stereo_kernel(const StereoJob* __restrict__ jobs, double hammingThreshold) {
  const StereoJob& job = jobs[blockIdx.y];
  const StereoSide& v0 = job.v0;
  const StereoSide& v1 = job.v1;
  const int k0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int n0 = v0.n_dev ? min(*v0.n_dev, v0.n) : v0.n;
  if (k0 >= n0) return;
  if (MOTION && !stereo_takes_part(job, k0, true)) {
    job.dist[k0] = hammingThreshold;
    job.k1[k0] = -1;
    job.hp[4 * k0 + 0] = job.hp[4 * k0 + 1] = job.hp[4 * k0 + 2] = job.hp[4 * k0 + 3] = 0.0;
    job.init[k0] = 0;
    job.quality[k0] = 0.0;
    return;
  }
  // poses: by value, or from the batch slot's record (the job table then stays valid from frame to frame)
  const double* __restrict__ C0 = v0.pose_slot ? v0.pose_slot->C_WC : v0.C;
  const double* __restrict__ C1 = v1.pose_slot ? v1.pose_slot->C_WC : v1.C;
  const double* __restrict__ p0 = v0.pose_slot ? v0.pose_slot->r_WC : v0.r;
  const double* __restrict__ p1 = v1.pose_slot ? v1.pose_slot->r_WC : v1.r;
  const V3 r0 = v3(p0[0], p0[1], p0[2]), r1 = v3(p1[0], p1[1], p1[2]);
  const bool valid0 = v0.bp_valid[k0] != 0;
  // running state of the reference's loop over k1
  double distances = hammingThreshold, quality = 0.0;
  V3 hps = v3(0.0, 0.0, 0.0);
  int k1_match = MOTION ? -1 : 0;
  bool initialisable = false;
  const int cnt = valid0 ? job.cand_count[k0] : 0;
  if (cnt > 0) {
    const double size0 = side_size(v0, k0);
    const V3 e0_W = normalized(matvec(C0, v3(v0.e_C[3 * k0], v0.e_C[3 * k0 + 1], v0.e_C[3 * k0 + 2])));
    auto visit = [&](int kc, double dc) {   // the body of the reference's loop for a k1 with dist < threshold
      if (!(dc < distances)) return;
      if (!v1.bp_valid[kc]) return;
      const double sigma = MOTION ? size0 / v0.f * 0.125 : fmax(size0 / v0.f, side_size(v1, kc) / v1.f) * 0.125;
      const V3 e1_W = normalized(matvec(C1, v3(v1.e_C[3 * kc], v1.e_C[3 * kc + 1], v1.e_C[3 * kc + 2])));
      if (MOTION && dot(e0_W, e1_W) < 0.5) return;
      bool isValid = false, isParallel = false;
      const V3 hp = triangulate_fast(r0, e0_W, r1, e1_W, cos(2.6 * sigma), cos(6.0 * sigma), isValid, isParallel);
      if (MOTION && !isValid) return;
      // T_WC.inverse() * hp_W with w = 1: C^T * p + (-(C^T * r)) * 1; z / w compared with 0.1
      const V3 t0 = matTvec(C0, r0), t1 = matTvec(C1, r1);
      const V3 c0 = matTvec(C0, hp), c1 = matTvec(C1, hp);
      const double z0 = c0.z + (-t0.z) * 1.0, z1 = c1.z + (-t1.z) * 1.0;
      const double minDepth = MOTION ? 0.2 : 0.1;
      if (z0 / 1.0 < minDepth) isValid = false;
      if (z1 / 1.0 < minDepth) isValid = false;
      if (dot(e0_W, e1_W) < 0.8) isValid = false;
      if (isValid) {
        distances = dc;
        hps = hp;
        k1_match = kc;
        initialisable = !isParallel;
        if (MOTION) quality = acos(dot(normalized(hp - r0), normalized(hp - r1)));
      }
    };
    if (cnt <= kCandPerQuery) {   // the listed candidates, in k1 order
      const unsigned* __restrict__ list = job.cand + (size_t)k0 * kCandPerQuery;
      unsigned last = 0;
      for (int i = 0; i < cnt; ++i) {
        last = next_candidate(list, cnt, last, i == 0);
        visit((int)(last >> kCandRowShift), (double)(last & ((1u << kCandRowShift) - 1u)));
      }
    } else {                      // the list overflowed (or no filter ran): every row, as the reference does
      const int n1 = stereo_rows(job);
      const StereoRowOf rowOf{job.v1_list, job.v1_owner_a, job.v1_owner_b};
      unsigned q[kDescWords];
      load_desc(v0.desc, k0, q);
      for (int i = 0; i < n1; ++i) {
        int kc;
        if (!rowOf(i, n1, kc)) continue;
        unsigned r[kDescWords];
        load_desc(v1.desc, kc, r);
        visit(kc, (double)hamming384(q, r));
      }
    }
  }
  job.dist[k0] = distances;
  job.k1[k0] = k1_match;
  job.hp[4 * k0 + 0] = hps.x / 1.0;
  job.hp[4 * k0 + 1] = hps.y / 1.0;
  job.hp[4 * k0 + 2] = hps.z / 1.0;
  job.hp[4 * k0 + 3] = (MOTION && k1_match < 0) ? 0.0 : 1.0;
  job.init[k0] = initialisable ? 1 : 0;
  if (MOTION) job.quality[k0] = quality;
}

// cand_counts: the jobs' candidate counters (one int per query), n_counts of them, cleared here; max1: upper bound of
// the rows (k1) of a job
void launch_match_stereo_jobs(const StereoJob* jobs, int nJobs, int max0, int max1, bool motion, double hammingThreshold,
                              int* cand_counts, size_t n_counts, cudaStream_t s) {
  if (nJobs <= 0 || max0 <= 0) return;
  const int thr = (int)ceil(hammingThreshold);
  const bool filter = max1 > 0 && max1 < kCandMaxRows && thr <= 385;
  // (counters of 0x7f7f7f7f = "overflow": every query then sweeps the rows itself)
  cudaMemsetAsync(cand_counts, filter ? 0 : 0x7f, n_counts * sizeof(int), s);
  if (filter) {
    const dim3 fgrid(filter_ctas_per_job(max0, max1, nJobs), nJobs);
    if (motion) stereo_filter_kernel<true><<<fgrid, 128, 0, s>>>(jobs, thr);
    else stereo_filter_kernel<false><<<fgrid, 128, 0, s>>>(jobs, thr);
  }
  const dim3 grid((max0 + 127) / 128, nJobs);
  if (motion) match_stereo_kernel<true><<<grid, 128, 0, s>>>(jobs, hammingThreshold);
  else match_stereo_kernel<false><<<grid, 128, 0, s>>>(jobs, hammingThreshold);
}

// ----------------------------------------------------------------------------------------------
// D. 3D-2D matching of uninitialised landmarks, Frontend::matchToMapByThreadUnitialised (Frontend.cpp:1830-1956)
// ----------------------------------------------------------------------------------------------
// Rows in landmark-major order. A row that beats the running best must pass, in this order:
// the epipolar-plane / divergence test unless the rays are (nearly) parallel (:1907-1917), triangulateFast valid
// (:1922-1927), >= 0.2 m from both centres (:1930-1936). Loop-closure rule (:1941-1944): an accepted row of the
// keypoint's previous landmark only counts and ends that landmark's rows for this keypoint. hps_W is overwritten
// only by non-parallel accepts (stale value kept otherwise, :1946-1951).
// Same two kernels as the 2D-2D matchers above: uninit_filter_kernel lists the rows below the threshold per keypoint,
// match_uninitialised_kernel resolves a keypoint per warp.
__device__ __forceinline__ bool uninit_takes_part(const UninitJob& job, int k) {
  // Frontend.cpp:1862-1872: a keypoint takes part when its back-projection is valid and it has no landmark yet
  if (job.use && job.use[k] == 0) return false;
  if (job.bp_valid && !job.bp_valid[k]) return false;
  if (job.owner14 && job.owner14[k] >= 0) return false;
  return true;
}
struct UninitRowOf {
  const int* rowList;
  const uint8_t* is3d;
  const int* owner;
  __device__ __forceinline__ bool operator()(int i, int hi, int& m) const {
    if (i >= hi) return false;
    m = rowList ? rowList[i] : i;
    return !(!rowList && is3d && is3d[owner[m]]);
  }
};
__device__ __forceinline__ int uninit_rows(const UninitJob& job) {
  // rows to sweep: all of the pool, or a precompacted ascending list (the rows of the non-3D landmarks)
  return job.row_list ? min(*job.n_list_dev, job.nm) : (job.nm_dev ? min(*job.nm_dev, job.nm) : job.nm);
}

__global__ void __launch_bounds__(128) uninit_filter_kernel(const UninitJob* __restrict__ jobs) {
  __shared__ int s_list[32];
  __shared__ int s_n;
  __shared__ uint2 s_queue[4][kFilterQueue];
  const UninitJob& job = jobs[blockIdx.y];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nk = job.nk_dev ? min(*job.nk_dev, job.nk) : job.nk;
  const int nm = uninit_rows(job);
  // rows: the compact copy of the listed rows' descriptors when there is one (sweep position = list position; the
  // resolve kernel maps it back through row_list), else the pool through the list / the is3d test
  const bool compact = job.row_list && job.row_desc;
  const UninitRowOf rowOf{compact ? nullptr : job.row_list, compact ? nullptr : job.is3d, job.owner};
  const uint8_t* __restrict__ rowDesc = compact ? job.row_desc : job.map;
  const int thr = (int)ceil(job.a.hammingThreshold);
  // work items of the job: (chunk of 32 keypoints) x (chunk of kFilterRows rows), dealt round-robin to the job's CTAs
  const int kChunks = (nk + 31) / 32, items = kChunks * ((nm + kFilterRows - 1) / kFilterRows);
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int kBase = (item % kChunks) * 32, lo = (item / kChunks) * kFilterRows;
    const int hi = min(nm, lo + kFilterRows);
    __syncthreads();
    if (warp == 0) {
      const int k = kBase + lane;
      const bool use = k < nk && uninit_takes_part(job, k);
      const unsigned bal = __ballot_sync(0xffffffffu, use);
      if (use) s_list[__popc(bal & ((1u << lane) - 1u))] = k;
      if (lane == 0) s_n = __popc(bal);
    }
    __syncthreads();
    const int nUse = s_n;
    for (int g0 = warp * kFilterQ; g0 < nUse; g0 += 4 * kFilterQ) {
      const int nq = min(kFilterQ, nUse - g0);
      unsigned q[kFilterQ][kDescWords];
      int out[kFilterQ];
#pragma unroll
      for (int j = 0; j < kFilterQ; ++j) {
        out[j] = s_list[g0 + min(j, nq - 1)];
        load_desc(job.queries, out[j], q[j]);
      }
      filter_sweep(job.queries, rowDesc, rowOf, lo, hi, q, nq, out, thr, job.cand, job.cand_count, s_queue[warp]);
    }
  }
}

// resolve: one thread per keypoint (see match_stereo_kernel)
__global__ void __launch_bounds__(128) match_uninitialised_kernel(const UninitJob* __restrict__ jobs) {
  const UninitJob& job = jobs[blockIdx.y];
  const UninitArgs& a = job.a;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int nk = job.nk_dev ? min(*job.nk_dev, job.nk) : job.nk;
  if (k >= nk) return;
  // running state of the reference's loop over the rows
  double best = a.hammingThreshold;
  int bestO = -1, skipOwner = -1;
  bool hpWritten = false;
  unsigned ctr = 0;
  V3 hps = v3(0.0, 0.0, 0.0);
  const int cnt = uninit_takes_part(job, k) ? job.cand_count[k] : 0;
  if (cnt > 0) {
    const int* __restrict__ owner = job.owner;
    const double* __restrict__ map_e_W = job.map_e_W;
    const double* __restrict__ map_r_W = job.map_r_W;
    V3 e1_W;
    if (job.kp_e_W) {
      e1_W = v3(job.kp_e_W[3 * k], job.kp_e_W[3 * k + 1], job.kp_e_W[3 * k + 2]);
    } else {   // T_WC1.C() * e1_C.normalized()
      e1_W = matvec(job.pose_slot ? job.pose_slot->C_WC : job.C_WC,
                    normalized(v3(job.e_C[3 * k], job.e_C[3 * k + 1], job.e_C[3 * k + 2])));
    }
    const double* __restrict__ rc = job.pose_slot ? job.pose_slot->r_WC : a.r_WC;
    const V3 rC = v3(rc[0], rc[1], rc[2]);
    const int prev = job.previousOwner ? job.previousOwner[k] : -1;
    const bool compact = job.row_list && job.row_desc;   // candidates are list positions (uninit_filter_kernel)
    auto visit = [&](int m, double dc) {   // the body of the reference's loop for a row with dist < threshold
      if (!(dc < best)) return;
      const int oc = owner[m];
      if (oc == skipOwner) return;  // "break" out of this landmark's descriptor loop happened earlier
      const V3 e0_W = v3(map_e_W[3 * m], map_e_W[3 * m + 1], map_e_W[3 * m + 2]);
      const V3 r0_W = v3(map_r_W[3 * m], map_r_W[3 * m + 1], map_r_W[3 * m + 2]);
      if (dot(e0_W, e1_W) < a.cos6Sigma) {
        const V3 et_W = normalized(rC - r0_W);
        const V3 n0_W = normalized(cross(e0_W, et_W));
        const V3 n1_W = normalized(cross(e1_W, et_W));
        if (dot(n0_W, n1_W) < a.cos6Sigma) return;                             // not in the epipolar plane
        if (dot(cross(e0_W, e1_W), normalized(n0_W + n0_W)) > 0.0) return;     // divergent rays ("n0 + n0" sic)
      }
      bool isValid = false, isParallel = false;
      const V3 hp = triangulate_fast(r0_W, e0_W, rC, e1_W, a.cos26Sigma, a.cos6Sigma, isValid, isParallel);
      if (!isValid) return;
      const V3 p_W = v3(hp.x / 1.0, hp.y / 1.0, hp.z / 1.0);
      if (norm(p_W - r0_W) < 0.2) isValid = false;
      if (norm(p_W - rC) < 0.2) isValid = false;
      if (!isValid) return;
      if (oc == prev) {               // loop-closure rule (:1941-1944)
        ++ctr;
        skipOwner = oc;
        return;
      }
      best = dc;
      bestO = oc;
      ++ctr;
      if (!isParallel) {
        hps = hp;
        hpWritten = true;
      }
    };
    if (cnt <= kCandPerQuery) {   // the listed candidates, in row order
      const unsigned* __restrict__ list = job.cand + (size_t)k * kCandPerQuery;
      unsigned last = 0;
      for (int i = 0; i < cnt; ++i) {
        last = next_candidate(list, cnt, last, i == 0);
        const int pos = (int)(last >> kCandRowShift);
        visit(compact ? job.row_list[pos] : pos, (double)(last & ((1u << kCandRowShift) - 1u)));
      }
    } else {                      // the list overflowed (or no filter ran): every row, as the reference does
      const int nm = uninit_rows(job);
      const UninitRowOf rowOf{job.row_list, job.is3d, job.owner};
      unsigned q[kDescWords];
      load_desc(job.queries, k, q);
      for (int i = 0; i < nm; ++i) {
        int m;
        if (!rowOf(i, nm, m)) continue;
        unsigned r[kDescWords];
        load_desc(job.map, m, r);
        visit(m, (double)hamming384(q, r));
      }
    }
  }
  job.dist[k] = best;
  job.best[k] = bestO;
  job.hp[4 * k + 0] = hps.x;
  job.hp[4 * k + 1] = hps.y;
  job.hp[4 * k + 2] = hps.z;
  job.hp[4 * k + 3] = hpWritten ? 1.0 : 0.0;
  job.hp_written[k] = hpWritten ? 1 : 0;
  if (job.ctr) job.ctr[k] = ctr;
  if (job.slice_ctrs && ctr) {
    // ctrs[t] of thread slice t (Frontend.cpp:1853-1855): segment = numKeypoints / numThreads, the last takes the rest
    const int segment = nk / job.num_slices;
    int t = segment == 0 ? job.num_slices - 1 : k / segment;
    if (t >= job.num_slices) t = job.num_slices - 1;
    atomicAdd(job.slice_ctrs + t, (unsigned long long)ctr);
  }
}

// ordered compaction of a slot's keypoints without a landmark (one CTA per slot)
__global__ void __launch_bounds__(1024) unmatched_lists_kernel(const UnmatchedArgs* __restrict__ args, int maxKp) {
  const UnmatchedArgs& a = args[blockIdx.x];
  __shared__ int s_warp[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nk = min(*a.nk_dev, maxKp);
  int* __restrict__ list = a.list + (size_t)blockIdx.x * maxKp;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int k0 = 0; k0 < nk; k0 += 1024) {
    const int k = k0 + tid;
    const bool un = k < nk && !(a.owner_a && a.owner_a[k] >= 0) && !(a.owner_b && a.owner_b[k] >= 0);
    const unsigned b = __ballot_sync(0xffffffffu, un);
    if (lane == 0) s_warp[warp] = __popc(b);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    if (un) list[before + __popc(b & ((1u << lane) - 1u))] = k;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += s_warp[w];
      s_base += t;
    }
    __syncthreads();
  }
  if (tid == 0) a.count[blockIdx.x] = s_base;
}

// One CTA per view: the live rows of a slot's keypoints / descriptors / back-projections / validity flags and their
// count into an exchange record (16-byte moves; every section of the record and of the arena is 16-byte aligned).
__global__ void __launch_bounds__(256) pack_views_kernel(const Slot* __restrict__ slots, const int* __restrict__ slotIndex,
                                                        int maxKp, PackLayout L, uint8_t* __restrict__ dst) {
  const Slot& slot = slots[slotIndex[blockIdx.x]];
  uint8_t* __restrict__ rec = dst + L.bytes * (size_t)blockIdx.x;
  const int n = min(slot.counters[2], maxKp);
  if (threadIdx.x == 0) *reinterpret_cast<int*>(rec + L.off_n) = n;
  auto copy = [&](const void* src, size_t off, size_t bytes) {
    const uint4* s4 = static_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(rec + off);
    const size_t q = (bytes + 15) / 16;
    for (size_t i = threadIdx.x; i < q; i += blockDim.x) d4[i] = s4[i];
  };
  copy(slot.kp_desc, L.off_kp, (size_t)n * sizeof(okc_keypoint));
  copy(slot.desc, L.off_desc, (size_t)n * 48);
  // (the back-projection arrays of a slot are only 8- / 1-byte aligned for an arbitrary max_keypoints)
  double* __restrict__ de = reinterpret_cast<double*>(rec + L.off_eC);
  for (int i = threadIdx.x; i < 3 * n; i += blockDim.x) de[i] = slot.e_C[i];
  for (int i = threadIdx.x; i < n; i += blockDim.x) rec[L.off_valid + i] = slot.bp_valid[i];
}

void launch_pack_views(const Slot* slots, const int* slot_index, int n, int max_kp, PackLayout layout, uint8_t* dst,
                       cudaStream_t s) {
  if (n > 0) pack_views_kernel<<<n, 256, 0, s>>>(slots, slot_index, max_kp, layout, dst);
}

void launch_unmatched_lists(const UnmatchedArgs* args, int n, int max_kp, cudaStream_t s) {
  if (n > 0 && max_kp > 0) unmatched_lists_kernel<<<n, 1024, 0, s>>>(args, max_kp);
}

// cand_counts: the jobs' candidate counters (one int per keypoint), n_counts of them, cleared here; maxRows: upper
// bound of the rows a job sweeps
void launch_match_uninit_jobs(const UninitJob* jobs, int nJobs, int maxNk, int maxRows, double hammingThreshold,
                              int* cand_counts, size_t n_counts, cudaStream_t s) {
  if (nJobs <= 0 || maxNk <= 0) return;
  const bool filter = maxRows > 0 && maxRows < kCandMaxRows && hammingThreshold <= 385.0;
  cudaMemsetAsync(cand_counts, filter ? 0 : 0x7f, n_counts * sizeof(int), s);   // (0x7f7f7f7f = "overflow": full sweeps)
  if (filter)
    uninit_filter_kernel<<<dim3(filter_ctas_per_job(maxNk, maxRows, nJobs), nJobs), 128, 0, s>>>(jobs);
  match_uninitialised_kernel<<<dim3((maxNk + 127) / 128, nJobs), 128, 0, s>>>(jobs);
}

}  // namespace okc
