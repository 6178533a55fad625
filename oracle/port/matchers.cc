// TEST INFRASTRUCTURE ONLY (oracle, "port" kind). CPU restatement in plain C++ doubles of the matching loops
// of okvis_frontend/src/Frontend.cpp, which cannot be compiled here (they need Eigen, glog, ceres, DBoW2,
// opengv; none is in this image). Parity of the FP64 gates is therefore UNPINNED by the reference itself:
// integer parts (Hamming distances, strict-'<' running minimum, iteration order) are exact restatements; the
// Eigen expressions are written out per SURVEY.md Appendix C (dot of 3-vectors = (x0*y0 + x1*y1) + x2*y2,
// normalized() = v / sqrt(v.v), computeInverseWithCheck = |det| > 1e-12).
//
// Built with the reference's flags (-O3 -mssse3, no FMA, no fast-math), see oracle/Makefile.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

// PORT_EIGEN_VARIANT (sensitivity builds, tests/test_eigen_order_sensitivity.py): the other evaluation orders an Eigen
// version / flag set could produce for the same expressions. bit 0: 3-vector dot as x0 y0 + (x1 y1 + x2 y2) (Eigen's
// non-vectorised unrolled redux); bit 1: normalized() as v * (1 / sqrt(v.v)); bit 2: 2x2 inverse by per-element
// division instead of a multiplication with 1 / det; bit 3: 3x3 * 3x1 rows as c0 x + (c1 y + c2 z).
#ifndef PORT_EIGEN_VARIANT
#define PORT_EIGEN_VARIANT 0
#endif

namespace {

// brisk::Hamming::PopcntofXORed(a, b, 3) (external/brisk/include/brisk/internal/hamming-inl.h:95-144):
// popcount of the XOR over 3 x 128 bits.
inline uint32_t hamming48(const uint8_t* a, const uint8_t* b) {
  uint64_t x[6], y[6];
  std::memcpy(x, a, 48);
  std::memcpy(y, b, 48);
  uint32_t d = 0;
  for (int i = 0; i < 6; ++i) d += static_cast<uint32_t>(__builtin_popcountll(x[i] ^ y[i]));
  return d;
}

struct V3 {
  double x, y, z;
};
inline V3 mk(double x, double y, double z) { return V3{x, y, z}; }
inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 scale(double s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline double dot(V3 a, V3 b) {
  return (PORT_EIGEN_VARIANT & 1) ? a.x * b.x + (a.y * b.y + a.z * b.z) : (a.x * b.x + a.y * b.y) + a.z * b.z;
}
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalized(V3 a) {
  const double z = dot(a, a);
  if (z > 0.0) {
    const double n = std::sqrt(z);
    if (PORT_EIGEN_VARIANT & 2) {
      const double r = 1.0 / n;
      return mk(a.x * r, a.y * r, a.z * r);
    }
    return mk(a.x / n, a.y / n, a.z / n);
  }
  return a;
}
inline double row3(double c0, double c1, double c2, V3 v) {
  return (PORT_EIGEN_VARIANT & 8) ? c0 * v.x + (c1 * v.y + c2 * v.z) : (c0 * v.x + c1 * v.y) + c2 * v.z;
}
inline V3 matvec(const double* C, V3 v) {
  return mk(row3(C[0], C[1], C[2], v), row3(C[3], C[4], C[5], v), row3(C[6], C[7], C[8], v));
}
inline V3 matTvec(const double* C, V3 v) {
  return mk(row3(C[0], C[3], C[6], v), row3(C[1], C[4], C[7], v), row3(C[2], C[5], C[8], v));
}

// triangulation::triangulateFast, okvis_frontend/src/stereo_triangulation.cpp:30-111
V3 triangulateFast(V3 p1, V3 e1, V3 p2, V3 e2, double sigma, bool& isValid, bool& isParallel) {
  isParallel = false;
  isValid = true;
  const V3 t12 = sub(p2, p1);
  const double b0 = dot(t12, e1), b1 = dot(t12, e2);
  const double A00 = dot(e1, e1), A10 = dot(e1, e2), A01 = -A10, A11 = -dot(e2, e2);
  const double det = A00 * A11 - A01 * A10;
  const bool invertible = std::fabs(det) > 1.0e-12;
  double l0 = 0.0, l1 = 0.0;
  if (invertible) {
    const double invdet = 1.0 / det;
    const bool byDiv = (PORT_EIGEN_VARIANT & 4) != 0;
    const double i00 = byDiv ? A11 / det : A11 * invdet, i01 = byDiv ? -A01 / det : -A01 * invdet,
                 i10 = byDiv ? -A10 / det : -A10 * invdet, i11 = byDiv ? A00 / det : A00 * invdet;
    l0 = i00 * b0 + i01 * b1;
    l1 = i10 * b0 + i11 * b1;
  }
  if (!invertible || l0 < 0.01 || l1 < 0.01) {  // :55-91 (both early-return blocks are the same code)
    isParallel = true;
    const V3 m = add(p1, scale(0.5, t12));
    const V3 midpoint = add(m, scale(40.0 * std::max(0.01, norm(t12)), add(e1, e2)));
    if (dot(e1, normalized(sub(midpoint, p1))) < std::cos(2.6 * sigma)) isValid = false;
    if (dot(e2, normalized(sub(midpoint, p2))) < std::cos(2.6 * sigma)) isValid = false;
    return midpoint;
  }
  const V3 xm = add(scale(l0, e1), p1);
  const V3 xn = add(scale(l1, e2), p2);
  const V3 s = add(xm, xn);
  const V3 midpoint = mk(s.x / 2.0, s.y / 2.0, s.z / 2.0);
  if (dot(e1, normalized(sub(midpoint, p1))) < std::cos(2.6 * sigma)) isValid = false;
  if (dot(e2, normalized(sub(midpoint, p2))) < std::cos(2.6 * sigma)) isValid = false;
  if (dot(normalized(sub(midpoint, p2)), normalized(sub(midpoint, p1))) > std::cos(6.0 * sigma)) isParallel = true;
  return midpoint;
}

}  // namespace

extern "C" {

uint32_t port_hamming(const uint8_t* a, const uint8_t* b) { return hamming48(a, b); }

// Plain best match, strict '<', first wins (Frontend::verifyRecognisedPlace, Frontend.cpp:316-345 inner loop).
// Multi-threaded over queries (nthreads) for the CPU baseline of the sweep.
void port_hamming_best(const uint8_t* queries, int nq, const uint8_t* map, int64_t nm, uint32_t threshold,
                       uint64_t indexOffset, uint64_t* best, int nthreads) {
  auto work = [&](int q0, int q1) {
    for (int q = q0; q < q1; ++q) {
      uint32_t bd = threshold;
      int64_t bi = -1;
      const uint8_t* qd = queries + 48 * static_cast<size_t>(q);
      for (int64_t m = 0; m < nm; ++m) {
        const uint32_t d = hamming48(qd, map + 48 * static_cast<size_t>(m));
        if (d < bd) {
          bd = d;
          bi = m;
        }
      }
      best[q] = bi < 0 ? ~0ull : (static_cast<uint64_t>(bd) << 32) | (static_cast<uint64_t>(bi) + indexOffset);
    }
  };
  if (nthreads <= 1) {
    work(0, nq);
    return;
  }
  std::vector<std::thread> th;
  const int per = (nq + nthreads - 1) / nthreads;
  for (int t = 0; t < nthreads; ++t) {
    const int a = t * per, b = std::min(nq, a + per);
    if (a < b) th.emplace_back(work, a, b);
  }
  for (auto& t : th) t.join();
}

// Frontend::matchToMapByThread for thread slice threadIdx of numThreads (Frontend.cpp:1744-1827).
// Landmarks are slots 0..L-1 in std::map (ascending id) order; rows of landmark l are
// map[rowStart[l] .. rowStart[l+1]). distances/lmIds must be initialised by the caller
// (distances = threshold, lmIds = -1), as Frontend.cpp:1511-1514 does.
void port_match_to_map_thread(int threadIdx, int numThreads, const uint8_t* ddata, const double* kp_xy,
                              const uint8_t* use_in, int numKeypoints, const uint8_t* map, const int32_t* rowStart,
                              int numLandmarks, const double* proj_xy, const uint8_t* is3d, double threshSq,
                              double* distances, int32_t* lmIds, uint64_t* ctrs, double* reprErrors) {
  ctrs[threadIdx] = 0;
  const size_t segment = static_cast<size_t>(numKeypoints) / numThreads;
  const size_t startK = segment * threadIdx;
  const size_t endK = threadIdx + 1 == numThreads ? static_cast<size_t>(numKeypoints) : startK + segment;
  for (int l = 0; l < numLandmarks; ++l) {
    if (is3d && !is3d[l]) continue;
    const double px = proj_xy[2 * l], py = proj_xy[2 * l + 1];
    for (size_t k = startK; k < endK; k++) {
      if (!use_in[k]) continue;
      const double rx = px - kp_xy[2 * k], ry = py - kp_xy[2 * k + 1];
      const double dd = rx * rx + ry * ry;
      if (dd > threshSq) continue;
      const uint8_t* descriptorK = ddata + k * 48;
      for (int r = rowStart[l]; r < rowStart[l + 1]; ++r) {
        const double dist = hamming48(descriptorK, map + static_cast<size_t>(r) * 48);
        if (dist < distances[k]) {
          distances[k] = dist;
          lmIds[k] = l;
          ctrs[threadIdx]++;
          reprErrors[threadIdx] += std::sqrt(dd);
        }
      }
    }
  }
  reprErrors[threadIdx] /= double(ctrs[threadIdx]);
}

// Frontend::matchStereo inner loops for one camera pair (Frontend.cpp:2259-2316).
struct PortStereoView {
  const uint8_t* desc;
  const double* size;
  const double* e_C;
  const uint8_t* bp_valid;
  int n;
  double C[9];
  double r[3];
  double f;
};
void port_match_stereo(const PortStereoView* v0, const PortStereoView* v1, double threshold, double* outDist,
                       int32_t* outK1, double* outHp, uint8_t* outInit) {
  const V3 r0 = mk(v0->r[0], v0->r[1], v0->r[2]), r1 = mk(v1->r[0], v1->r[1], v1->r[2]);
  for (int k0 = 0; k0 < v0->n; ++k0) {
    double distances = threshold;
    bool initialisable = false;
    V3 hps = mk(0, 0, 0);
    int k1_match = 0;
    for (int k1 = 0; k1 < v1->n; ++k1) {
      const double dist = hamming48(v0->desc + 48 * static_cast<size_t>(k0), v1->desc + 48 * static_cast<size_t>(k1));
      if (dist < distances) {
        const double sigma = std::max(v0->size[k0] / v0->f, v1->size[k1] / v1->f) * 0.125;
        bool isValid = false, isParallel = false;
        if (!v0->bp_valid[k0]) continue;
        if (!v1->bp_valid[k1]) continue;
        const V3 e0_W = normalized(matvec(v0->C, mk(v0->e_C[3 * k0], v0->e_C[3 * k0 + 1], v0->e_C[3 * k0 + 2])));
        const V3 e1_W = normalized(matvec(v1->C, mk(v1->e_C[3 * k1], v1->e_C[3 * k1 + 1], v1->e_C[3 * k1 + 2])));
        const V3 hp = triangulateFast(r0, e0_W, r1, e1_W, sigma, isValid, isParallel);
        // hp_C = T_WC.inverse() * hp_W; T^-1 = [C^T, -C^T r]
        const V3 t0 = matTvec(v0->C, r0), t1 = matTvec(v1->C, r1);
        const V3 c0 = matTvec(v0->C, hp), c1 = matTvec(v1->C, hp);
        const double z0 = c0.z + (-t0.z) * 1.0, z1 = c1.z + (-t1.z) * 1.0;
        if (z0 / 1.0 < 0.1) isValid = false;
        if (z1 / 1.0 < 0.1) isValid = false;
        if (dot(e0_W, e1_W) < 0.8) isValid = false;
        if (isValid) {
          distances = dist;
          hps = hp;
          k1_match = k1;
          initialisable = !isParallel;
        }
      }
    }
    outDist[k0] = distances;
    outK1[k0] = k1_match;
    outHp[4 * k0 + 0] = hps.x / 1.0;
    outHp[4 * k0 + 1] = hps.y / 1.0;
    outHp[4 * k0 + 2] = hps.z / 1.0;
    outHp[4 * k0 + 3] = 1.0;
    outInit[k0] = initialisable ? 1 : 0;
  }
}

// Frontend::matchToMapByThreadUnitialised for thread slice threadIdx of numThreads (Frontend.cpp:1830-1956).
// The caller passes only the landmarks this call iterates over (not is3d, and inside the loop-closure set when one
// is given, :1876-1886) as slots 0..L-1 in std::map order; rows of slot l are map[rowStart[l] .. rowStart[l+1])
// with their rays e_W / origins r_W (LandmarkToMatch::e_W, r_W columns). kp_e_W[k] = T_WC1.C() * e1_C.normalized()
// (:1862-1863); use[k] as computed at :1860-1870; previousOwner[k] = slot of multiFrame->landmarkId(im,k) or -1.
// distances / lmIds / hps_W (xyzw, zero-initialised) are in/out like the reference's vectors.
void port_match_uninit_thread(int threadIdx, int numThreads, const uint8_t* ddata, const double* kp_e_W,
                              const uint8_t* use, const int32_t* previousOwner, int numKeypoints, const uint8_t* map,
                              const int32_t* rowStart, int numLandmarks, const double* map_e_W,
                              const double* map_r_W, const double* r_WC, double focalLength, double* distances,
                              int32_t* lmIds, double* hps_W, uint8_t* hpsWritten, uint64_t* ctrs) {
  ctrs[threadIdx] = 0;
  const size_t segment = static_cast<size_t>(numKeypoints) / numThreads;
  const size_t startK = segment * threadIdx;
  const size_t endK = threadIdx + 1 == numThreads ? static_cast<size_t>(numKeypoints) : startK + segment;
  const double sigma = 1.0 / focalLength;
  const double cos6Sigma = std::cos(6.0 * sigma);
  const V3 rC = mk(r_WC[0], r_WC[1], r_WC[2]);
  for (int l = 0; l < numLandmarks; ++l) {
    for (size_t k = startK; k < endK; k++) {
      if (!use[k]) continue;
      const V3 e1_W = mk(kp_e_W[3 * k], kp_e_W[3 * k + 1], kp_e_W[3 * k + 2]);
      const uint8_t* descriptorK = ddata + k * 48;
      for (int r = rowStart[l]; r < rowStart[l + 1]; ++r) {
        const double dist = hamming48(descriptorK, map + static_cast<size_t>(r) * 48);
        if (dist < distances[k]) {
          const V3 e0_W = mk(map_e_W[3 * r], map_e_W[3 * r + 1], map_e_W[3 * r + 2]);
          const V3 r0_W = mk(map_r_W[3 * r], map_r_W[3 * r + 1], map_r_W[3 * r + 2]);
          if (dot(e0_W, e1_W) < cos6Sigma) {
            const V3 et_W = normalized(sub(rC, r0_W));
            const V3 n0_W = normalized(cross(e0_W, et_W));
            const V3 n1_W = normalized(cross(e1_W, et_W));
            if (dot(n0_W, n1_W) < cos6Sigma) continue;
            if (dot(cross(e0_W, e1_W), normalized(add(n0_W, n0_W))) > 0.0) continue;  // "n0_W + n0_W" sic (:1914)
          }
          bool isValid = false, isParallel = false;
          const V3 hp = triangulateFast(r0_W, e0_W, rC, e1_W, sigma, isValid, isParallel);
          if (!isValid) continue;
          const V3 p_W = mk(hp.x / 1.0, hp.y / 1.0, hp.z / 1.0);
          if (norm(sub(p_W, r0_W)) < 0.2) isValid = false;
          if (norm(sub(p_W, rC)) < 0.2) isValid = false;
          if (!isValid) continue;
          if (previousOwner && l == previousOwner[k]) {
            ctrs[threadIdx]++;
            break;
          }
          distances[k] = dist;
          lmIds[k] = l;
          ctrs[threadIdx]++;
          if (!isParallel) {
            hps_W[4 * k + 0] = hp.x;
            hps_W[4 * k + 1] = hp.y;
            hps_W[4 * k + 2] = hp.z;
            hps_W[4 * k + 3] = 1.0;
            hpsWritten[k] = 1;
          }
        }
      }
    }
  }
}

// Worker lambda of Frontend::matchMotionStereo for all k0 (Frontend.cpp:2053-2137), up to (not including) the
// reprojection check of :2140-2147. use0[k0] folds the estimator-state tests (:2056-2064, 2082-2084).
void port_match_motion_stereo(const PortStereoView* v0, const uint8_t* use0, const PortStereoView* v1, double threshold,
                              double* outDist, int32_t* outK1, double* outHp, uint8_t* outInit, double* outQuality) {
  const V3 r0 = mk(v0->r[0], v0->r[1], v0->r[2]), r1 = mk(v1->r[0], v1->r[1], v1->r[2]);
  for (int k0 = 0; k0 < v0->n; ++k0) {
    uint32_t distances = static_cast<uint32_t>(threshold);
    bool initialisable = false;
    double quality = 0.0;
    V3 hps = mk(0, 0, 0);
    double hw = 0.0;
    int k1_max = -1;
    if (use0[k0] && v0->bp_valid[k0]) {
      const V3 e0_W = normalized(matvec(v0->C, mk(v0->e_C[3 * k0], v0->e_C[3 * k0 + 1], v0->e_C[3 * k0 + 2])));
      const double sigma = v0->size[k0] / v0->f * 0.125;
      for (int kk = 0; kk < v1->n; ++kk) {
        const uint32_t dist = hamming48(v0->desc + 48 * static_cast<size_t>(k0), v1->desc + 48 * static_cast<size_t>(kk));
        if (dist < distances) {
          bool isValid = false, isParallel = false;
          if (!v1->bp_valid[kk]) continue;
          const V3 e1_W = normalized(matvec(v1->C, mk(v1->e_C[3 * kk], v1->e_C[3 * kk + 1], v1->e_C[3 * kk + 2])));
          if (dot(e0_W, e1_W) < 0.5) continue;
          const V3 hp = triangulateFast(r0, e0_W, r1, e1_W, sigma, isValid, isParallel);
          if (!isValid) continue;
          const V3 t0 = matTvec(v0->C, r0), t1 = matTvec(v1->C, r1);
          const V3 c0 = matTvec(v0->C, hp), c1 = matTvec(v1->C, hp);
          const double z0 = c0.z + (-t0.z) * 1.0, z1 = c1.z + (-t1.z) * 1.0;
          if (dot(e0_W, e1_W) < 0.8) isValid = false;
          if (z0 / 1.0 < 0.2) isValid = false;
          if (z1 / 1.0 < 0.2) isValid = false;
          if (isValid) {
            k1_max = kk;
            distances = dist;
            quality = std::acos(dot(normalized(sub(hp, r0)), normalized(sub(hp, r1))));
            hps = hp;
            hw = 1.0;
            initialisable = !isParallel;
          }
        }
      }
    }
    outDist[k0] = static_cast<double>(distances);
    outK1[k0] = k1_max;
    outHp[4 * k0 + 0] = hps.x;
    outHp[4 * k0 + 1] = hps.y;
    outHp[4 * k0 + 2] = hps.z;
    outHp[4 * k0 + 3] = hw;
    outInit[k0] = initialisable ? 1 : 0;
    outQuality[k0] = quality;
  }
}

}  // extern "C"
