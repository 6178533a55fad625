// TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the camera geometry next to the hot path.
//   port_back_project      - okvis::Frame::computeBackProjections (okvis_cv/include/okvis/implementation/Frame.hpp:158-173)
//                            through PinholeCamera::backProject (cameras/implementation/PinholeCamera.hpp:566-585) and
//                            the distortion models' undistort / distort
//                            (RadialTangentialDistortion.hpp:70-112, 194-233; EquidistantDistortion.hpp:87-189, 299-332)
//   port_project_landmarks - the FoV test of Frontend::matchToMap (okvis_frontend/src/Frontend.cpp:1376-1407)
// okvis_cv needs Eigen, which this image does not have, so the 2x2 / 3x3 Eigen expressions are written out with plain
// doubles (small fixed-size Eigen products are evaluated coefficient by coefficient; Matrix2d::inverse() is the
// adjugate scaled by 1/det). PARITY UNPINNED by upstream: the reference has no golden values for these functions.
#include <algorithm>
#include <cmath>
#include <cstdint>

namespace {

struct Pinhole {  // layout of okc_pinhole (include/okvis_cuda.h)
  int width, height;
  double fu, fv, cu, cv;
  int distortion;  // 0 none, 1 radial-tangential, 2 equidistant
  double k[8];
};

struct KeyPoint { float x, y, size, angle, response; int octave, class_id; };

typedef double M22[2][2];

bool distort(const Pinhole& c, const double u[2], double out[2], M22* Jp) {
  if (c.distortion == 1) {
    const double k1_ = c.k[0], k2_ = c.k[1], p1_ = c.k[2], p2_ = c.k[3];
    const double u0 = u[0], u1 = u[1];
    const double mx_u = u0 * u0;
    const double my_u = u1 * u1;
    const double mxy_u = u0 * u1;
    const double rho_u = mx_u + my_u;
    const double rad_dist_u = k1_ * rho_u + k2_ * rho_u * rho_u;
    out[0] = u0 + u0 * rad_dist_u + 2.0 * p1_ * mxy_u + p2_ * (rho_u + 2.0 * mx_u);
    out[1] = u1 + u1 * rad_dist_u + 2.0 * p2_ * mxy_u + p1_ * (rho_u + 2.0 * my_u);
    if (Jp) {
      M22& J = *Jp;
      J[0][0] = 1 + rad_dist_u + k1_ * 2.0 * mx_u + k2_ * rho_u * 4 * mx_u + 2.0 * p1_ * u1 + 6 * p2_ * u0;
      J[1][0] = k1_ * 2.0 * u0 * u1 + k2_ * 4 * rho_u * u0 * u1 + p1_ * 2.0 * u0 + 2.0 * p2_ * u1;
      J[0][1] = J[1][0];
      J[1][1] = 1 + rad_dist_u + k1_ * 2.0 * my_u + k2_ * rho_u * 4 * my_u + 6 * p1_ * u1 + 2.0 * p2_ * u0;
    }
    return true;
  }
  if (c.distortion == 2) {
    const double k1_ = c.k[0], k2_ = c.k[1], k3_ = c.k[2], k4_ = c.k[3];
    const double u0 = u[0], u1 = u[1];
    const double r = std::sqrt(u0 * u0 + u1 * u1);
    const double theta = std::atan(r);
    const double theta2 = theta * theta;
    const double theta4 = theta2 * theta2;
    const double theta6 = theta4 * theta2;
    const double theta8 = theta4 * theta4;
    const double thetad = theta * (1.0 + k1_ * theta2 + k2_ * theta4 + k3_ * theta6 + k4_ * theta8);
    const double scaling = (r > 1e-8) ? thetad / r : 1.0;
    out[0] = scaling * u0;
    out[1] = scaling * u1;
    if (Jp) {
      M22& J = *Jp;
      if (r > 1e-8) {
        double t2 = u0 * u0;
        double t3 = u1 * u1;
        double t4 = t2 + t3;
        double t6 = std::atan(std::sqrt(t4));
        double t7 = t6 * t6;
        double t8 = 1.0 / std::sqrt(t4);
        double t9 = t7 * t7;
        double t11 = 1.0 / ((t2 + t3) + 1.0);
        double t17 = (((k1_ * t7 + k2_ * t9) + k3_ * t7 * t9) + k4_ * (t9 * t9)) + 1.0;
        double t18 = 1.0 / t4;
        double t19 = 1.0 / std::sqrt(t4 * t4 * t4);
        double t20 = t6 * t8 * t17;
        double t25 = ((k2_ * t6 * t7 * t8 * t11 * u1 * 4.0 + k3_ * t6 * t8 * t9 * t11 * u1 * 6.0) +
                      k4_ * t6 * t7 * t8 * t9 * t11 * u1 * 8.0) +
                     k1_ * t6 * t8 * t11 * u1 * 2.0;
        t4 = ((k2_ * t6 * t7 * t8 * t11 * u0 * 4.0 + k3_ * t6 * t8 * t9 * t11 * u0 * 6.0) +
              k4_ * t6 * t7 * t8 * t9 * t11 * u0 * 8.0) +
             k1_ * t6 * t8 * t11 * u0 * 2.0;
        t7 = t11 * t17 * t18 * u0 * u1;
        J[0][1] = (t7 + t6 * t8 * t25 * u0) - t6 * t17 * t19 * u0 * u1;
        J[1][1] = ((t20 - t3 * t6 * t17 * t19) + t3 * t11 * t17 * t18) + t6 * t8 * t25 * u1;
        J[0][0] = ((t20 - t2 * t6 * t17 * t19) + t2 * t11 * t17 * t18) + t6 * t8 * t4 * u0;
        J[1][0] = (t7 + t6 * t8 * t4 * u1) - t6 * t17 * t19 * u0 * u1;
      } else {
        J[0][0] = J[1][1] = 1.0;
        J[0][1] = J[1][0] = 0.0;
      }
    }
    return true;
  }
  out[0] = u[0];
  out[1] = u[1];
  if (Jp) {
    (*Jp)[0][0] = (*Jp)[1][1] = 1.0;
    (*Jp)[0][1] = (*Jp)[1][0] = 0.0;
  }
  return true;
}

bool undistort(const Pinhole& c, const double y[2], double x[2]) {
  if (c.distortion == 0) {
    x[0] = y[0];
    x[1] = y[1];
    return true;
  }
  double x_bar[2] = {y[0], y[1]};
  const int n = c.distortion == 2 ? 20 : 5;
  bool success = false;
  for (int i = 0; i < n; i++) {
    double x_tmp[2];
    M22 E;
    distort(c, x_bar, x_tmp, &E);
    const double e[2] = {y[0] - x_tmp[0], y[1] - x_tmp[1]};
    M22 E2, inv, M;
    for (int r = 0; r < 2; ++r)
      for (int q = 0; q < 2; ++q) E2[r][q] = E[0][r] * E[0][q] + E[1][r] * E[1][q];   // E^T * E
    const double invdet = 1.0 / (E2[0][0] * E2[1][1] - E2[1][0] * E2[0][1]);
    inv[0][0] = E2[1][1] * invdet;
    inv[1][0] = -E2[1][0] * invdet;
    inv[0][1] = -E2[0][1] * invdet;
    inv[1][1] = E2[0][0] * invdet;
    for (int r = 0; r < 2; ++r)
      for (int q = 0; q < 2; ++q) M[r][q] = inv[r][0] * E[q][0] + inv[r][1] * E[q][1];  // inv * E^T
    for (int r = 0; r < 2; ++r) x_bar[r] += M[r][0] * e[0] + M[r][1] * e[1];
    const double chi2 = e[0] * e[0] + e[1] * e[1];
    if (chi2 < 1e-6) success = true;
    if (chi2 < 1e-15) {
      success = true;
      break;
    }
  }
  x[0] = x_bar[0];
  x[1] = x_bar[1];
  return success;
}

}  // namespace

extern "C" {

int port_back_project(const Pinhole* cam, const KeyPoint* kps, int n, double* e_C, uint8_t* valid) {
  int ctr = 0;
  const double one_over_fu = 1.0 / cam->fu, one_over_fv = 1.0 / cam->fv;
  for (int k = 0; k < n; ++k) {
    const double p2[2] = {((double)kps[k].x - cam->cu) * one_over_fu, ((double)kps[k].y - cam->cv) * one_over_fv};
    double und[2];
    const bool ok = undistort(*cam, p2, und);
    e_C[3 * k] = und[0];
    e_C[3 * k + 1] = und[1];
    e_C[3 * k + 2] = 1.0;
    valid[k] = ok;
    ctr += ok;
  }
  return ctr;
}

// status: 0 Successful, 1 OutsideImage, 2 Masked, 3 Behind, 4 Invalid (cameras::ProjectionStatus)
void port_project_landmarks(const Pinhole* cam, const double* T_CW, const double* hp_W, int n, double reprThreshold,
                            double* proj_xy, uint8_t* status, uint8_t* visible) {
  const double maxU = cam->width + reprThreshold, maxV = cam->height + reprThreshold;
  for (int l = 0; l < n; ++l) {
    const double* h = hp_W + 4 * l;
    double p[3];
    for (int i = 0; i < 3; ++i)
      p[i] = ((T_CW[3 * i] * h[0] + T_CW[3 * i + 1] * h[1]) + T_CW[3 * i + 2] * h[2]) + T_CW[9 + i] * h[3];
    if (h[3] < 0)
      for (int i = 0; i < 3; ++i) p[i] = -p[i];
    double kp[2] = {0.0, 0.0};
    int st;
    if (std::fabs(p[2]) < 1.0e-12) {
      st = 4;
    } else {
      const double rz = 1.0 / p[2];
      const double und[2] = {p[0] * rz, p[1] * rz};
      double dist[2];
      distort(*cam, und, dist, nullptr);
      kp[0] = cam->fu * dist[0] + cam->cu;
      kp[1] = cam->fv * dist[1] + cam->cv;
      if (kp[0] < 0.0 || kp[1] < 0.0 || kp[0] >= cam->width || kp[1] >= cam->height) st = 1;
      else st = p[2] > 0.0 ? 0 : 3;
    }
    bool vis = !(st == 4 || st == 3);
    if (vis && (kp[0] < -reprThreshold || kp[1] < -reprThreshold || kp[0] > maxU || kp[1] > maxV)) vis = false;
    proj_xy[2 * l] = kp[0];
    proj_xy[2 * l + 1] = kp[1];
    status[l] = (uint8_t)st;
    visible[l] = vis;
  }
}

}  // extern "C"

// Keyframe-overlap masks of Frontend::doWeNeedANewKeyframe (okvis_frontend/src/Frontend.cpp:1199-1231, 1255-1278):
// cv::circle(mask, keypoint.pt * 0.1, int(radius), 255, cv::FILLED) restated as the integer disc OpenCV draws
// (pinned against cv2 4.13 in tests/golden/overlap_goldens.npz), cv::bitwise_and / bitwise_or / countNonZero.
#include <cstring>
#include <vector>
extern "C" void port_keyframe_overlap(int n_views, const int* width, const int* height, const int64_t* kp_offset,
                                      const KeyPoint* kps, const uint8_t* matched, int32_t* inter, int32_t* uni) {
  for (int v = 0; v < n_views; ++v) {
    const int rows = height[v] / 10, cols = width[v] / 10;
    std::vector<uint8_t> det((size_t)rows * cols, 0), mat((size_t)rows * cols, 0);
    const double radius = double(rows < cols ? rows : cols) * 0.09;
    const int r = int(radius);
    for (int64_t k = kp_offset[v]; k < kp_offset[v + 1]; ++k) {
      const float fx = float(double(kps[k].x) * 0.1), fy = float(double(kps[k].y) * 0.1);   // Point2f * double
      const int cx = (int)lrintf(fx), cy = (int)lrintf(fy);                                 // cvRound
      for (int dy = -r; dy <= r; ++dy) {
        const int y = cy + dy;
        if (y < 0 || y >= rows) continue;
        int hw = 0;
        while ((hw + 1) * (hw + 1) + dy * dy <= r * r) ++hw;
        for (int x = cx - hw; x <= cx + hw; ++x) {
          if (x < 0 || x >= cols) continue;
          det[(size_t)y * cols + x] = 255;
          if (matched[k]) mat[(size_t)y * cols + x] = 255;
        }
      }
    }
    int ci = 0, cu = 0;
    for (size_t i = 0; i < det.size(); ++i) {
      ci += (det[i] & mat[i]) != 0;
      cu += (det[i] | mat[i]) != 0;
    }
    inter[v] = ci;
    uni[v] = cu;
  }
}

// Landmark preparation of Frontend::matchToMap, second half (okvis_frontend/src/Frontend.cpp:1409-1500): restated
// with plain doubles in the reference's statement order. Eigen pieces: normalized() = v / sqrt(v.v) (v.v > 0),
// 3-vector sums (x0 y0 + x1 y1) + x2 y2.
#ifndef PORT_EIGEN_VARIANT
#define PORT_EIGEN_VARIANT 0   // see matchers.cc
#endif
namespace {
struct V3 { double x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 mul(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(V3 a, V3 b) {
  return (PORT_EIGEN_VARIANT & 1) ? a.x * b.x + (a.y * b.y + a.z * b.z) : (a.x * b.x + a.y * b.y) + a.z * b.z;
}
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalized(V3 a) {
  const double n2 = dot(a, a);
  if (n2 > 0.0) {
    const double n = std::sqrt(n2);
    if (PORT_EIGEN_VARIANT & 2) {
      const double r = 1.0 / n;
      return {a.x * r, a.y * r, a.z * r};
    }
    return {a.x / n, a.y / n, a.z / n};
  }
  return a;
}
}  // namespace

extern "C" int port_prepare_landmarks(int n_views, const double* C_WC, const double* r_WC, const int64_t* row0,
                                      const uint8_t* desc, const double* e_C, int L, const double* hp_W,
                                      const double* quality, const uint8_t* visible, const int64_t* obs_offset,
                                      const int32_t* obs_view, const int32_t* obs_kp, const double* r_WC1,
                                      double focalLength, int loopClosure, uint8_t* is3dOut, int32_t* rowsOut,
                                      uint8_t* pool, int32_t* owner, double* e_W, double* r_Wout, uint8_t* filled,
                                      int32_t* row_obs) {
  (void)n_views;
  int nm = 0;
  for (int l = 0; l < L; ++l) {
    is3dOut[l] = 0;
    rowsOut[l] = 0;
    if (!visible[l]) continue;
    bool is3d = false;
    const double* hp = hp_W + 4 * l;
    const V3 p_W = {hp[0] / hp[3], hp[1] / hp[3], hp[2] / hp[3]};
    const V3 r_W = sub(p_W, V3{r_WC1[0], r_WC1[1], r_WC1[2]});
    const V3 e = normalized(r_W);
    const double r = std::max(0.01, norm(r_W));
    const double q = quality[l];
    double bestScores[3] = {1.0, 1.0, 1.0};
    int slot[3] = {-1, -1, -1};
    size_t o = 0;
    for (int64_t i = obs_offset[l]; i < obs_offset[l + 1]; ++i) {
      const int v = obs_view[i];
      const V3 r_old = {r_WC[3 * v], r_WC[3 * v + 1], r_WC[3 * v + 2]};
      const V3 r_W_old = sub(p_W, r_old);
      if (!is3d) {
        const V3 r_close_W = sub(r_W, mul(0.2 / focalLength / q, r_W_old));
        const double cosA = dot(normalized(r_W), normalized(r_close_W));
        if (cosA > std::cos(10.0 / focalLength)) is3d = true;
      }
      const double cosViewpointChange = dot(e, normalized(r_W_old));
      if (cosViewpointChange < std::cos(0.6) && !loopClosure) continue;
      const double scaleChange = std::fabs(r - norm(r_W_old)) / r;
      if ((scaleChange > 0.5) && !loopClosure) continue;
      const double score = 0.5 * (std::acos(cosViewpointChange) / 0.6 + scaleChange / 0.5);
      double worstScore = 0.0;
      size_t worstIdx = 0;
      for (size_t n = 0; n < 3; ++n) {
        if (bestScores[n] > worstScore) {
          worstScore = bestScores[n];
          worstIdx = n;
        }
      }
      if (score < bestScores[worstIdx]) {
        slot[worstIdx] = int(i - obs_offset[l]);
        o = std::max(o, worstIdx);
        bestScores[worstIdx] = score;
      }
    }
    is3dOut[l] = is3d;
    rowsOut[l] = int(o + 1);
    for (size_t n = 0; n <= o; ++n, ++nm) {
      owner[nm] = l;
      row_obs[nm] = slot[n];
      filled[nm] = slot[n] >= 0;
      if (slot[n] < 0) {
        std::memset(pool + (size_t)nm * 48, 0, 48);
        for (int c = 0; c < 3; ++c) e_W[3 * nm + c] = r_Wout[3 * nm + c] = 0.0;
        continue;
      }
      const int64_t i = obs_offset[l] + slot[n];
      const int v = obs_view[i];
      const int64_t src = row0[v] + obs_kp[i];
      std::memcpy(pool + (size_t)nm * 48, desc + (size_t)src * 48, 48);
      const V3 en = normalized(V3{e_C[3 * src], e_C[3 * src + 1], e_C[3 * src + 2]});
      const double* C = C_WC + 9 * v;
      for (int c = 0; c < 3; ++c) {
        e_W[3 * nm + c] = (PORT_EIGEN_VARIANT & 8) ? C[3 * c] * en.x + (C[3 * c + 1] * en.y + C[3 * c + 2] * en.z)
                                                   : (C[3 * c] * en.x + C[3 * c + 1] * en.y) + C[3 * c + 2] * en.z;
        r_Wout[3 * nm + c] = r_WC[3 * v + c];
      }
    }
  }
  return nm;
}

// DBoW2::TemplatedVocabulary<FBrisk>::transform of one feature (external/DBoW2/include/DBoW2/TemplatedVocabulary.h:
// 1216-1257) over a flattened tree: children[child_offset[id] .. child_offset[id+1]) in load order.
extern "C" void port_vocabulary_transform(const int32_t* child_offset, const int32_t* children, const uint8_t* node_desc,
                                          const double* node_weight, const int32_t* node_word, const uint8_t* features,
                                          int n, int m_L, int levelsup, uint32_t* word_id, double* weight,
                                          uint32_t* nid) {
  auto distance = [](const uint8_t* a, const uint8_t* b) {
    int d = 0;
    for (int i = 0; i < 48; ++i) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
    return double(d);
  };
  for (int f = 0; f < n; ++f) {
    const uint8_t* feature = features + (size_t)f * 48;
    const int nid_level = m_L - levelsup;
    nid[f] = 0;
    int final_id = 0;
    int current_level = 0;
    do {
      ++current_level;
      const int32_t* nodes = children + child_offset[final_id];
      const int count = child_offset[final_id + 1] - child_offset[final_id];
      final_id = nodes[0];
      double best_d = distance(feature, node_desc + (size_t)final_id * 48);
      for (int c = 1; c < count; ++c) {
        const int id = nodes[c];
        const double d = distance(feature, node_desc + (size_t)id * 48);
        if (d < best_d) {
          best_d = d;
          final_id = id;
        }
      }
      if (current_level == nid_level) nid[f] = (uint32_t)final_id;
    } while (child_offset[final_id + 1] > child_offset[final_id]);
    word_id[f] = (uint32_t)node_word[final_id];
    weight[f] = node_weight[final_id];
  }
}
