// Camera geometry next to the hot path (SURVEY.md section 8(f), rows f2 and the first half of f1).
//
//  back_project_kernel     = okvis::Frame::computeBackProjections (okvis_cv/include/okvis/implementation/Frame.hpp:158-173)
//                            -> PinholeCamera<D>::backProject (cameras/implementation/PinholeCamera.hpp:566-585)
//                            -> D::undistort, Gauss-Newton on D::distort
//                               (RadialTangentialDistortion.hpp:70-112, 194-233; EquidistantDistortion.hpp:87-160, 299-332)
//  project_landmarks_kernel = the field-of-view test at the top of the landmark loop of Frontend::matchToMap
//                            (okvis_frontend/src/Frontend.cpp:1376-1407): hp_C = T_CW * hp_W
//                            (kinematics/implementation/Transformation.hpp:251-258), projectHomogeneous
//                            (PinholeCamera.hpp:249-283, 485-496), the +-reprThreshold window.
//
// FP64, one thread per keypoint / landmark, compiled with --fmad=false like everything else: every operation is
// rounded separately, as in the reference's x86-64 build without FMA. The Eigen expressions are written out in the
// order Eigen evaluates them for fixed 2x2 / 3x3 sizes (inverse = adjugate * (1 / det); products coefficient by
// coefficient). Eigen itself is not in this image, so this order is restated, not pinned.
#include "common.cuh"
#include <cstring>
#include <new>
#include <vector>

namespace okc {

struct Mat2 { double a, b, c, d; };  // [a b; c d]

// D::distort with the 2x2 Jacobian wrt. the undistorted point. Returns false where the reference's distort does.
__host__ __device__ inline bool distort_point(const okc_pinhole& cam, double u0, double u1, double* d0, double* d1,
                                              Mat2* J) {
  if (cam.distortion == OKC_DISTORTION_RADTAN) {
    const double k1 = cam.k[0], k2 = cam.k[1], p1 = cam.k[2], p2 = cam.k[3];
    const double mx_u = u0 * u0, my_u = u1 * u1, mxy_u = u0 * u1;
    const double rho_u = mx_u + my_u;
    const double rad_dist_u = k1 * rho_u + k2 * rho_u * rho_u;
    *d0 = u0 + u0 * rad_dist_u + 2.0 * p1 * mxy_u + p2 * (rho_u + 2.0 * mx_u);
    *d1 = u1 + u1 * rad_dist_u + 2.0 * p2 * mxy_u + p1 * (rho_u + 2.0 * my_u);
    if (J) {
      J->a = 1 + rad_dist_u + k1 * 2.0 * mx_u + k2 * rho_u * 4 * mx_u + 2.0 * p1 * u1 + 6 * p2 * u0;
      J->c = k1 * 2.0 * u0 * u1 + k2 * 4 * rho_u * u0 * u1 + p1 * 2.0 * u0 + 2.0 * p2 * u1;
      J->b = J->c;
      J->d = 1 + rad_dist_u + k1 * 2.0 * my_u + k2 * rho_u * 4 * my_u + 6 * p1 * u1 + 2.0 * p2 * u0;
    }
    return true;
  }
  if (cam.distortion == OKC_DISTORTION_EQUIDISTANT) {
    const double k1 = cam.k[0], k2 = cam.k[1], k3 = cam.k[2], k4 = cam.k[3];
    const double r = sqrt(u0 * u0 + u1 * u1);
    const double theta = atan(r);
    const double theta2 = theta * theta;
    const double theta4 = theta2 * theta2;
    const double theta6 = theta4 * theta2;
    const double theta8 = theta4 * theta4;
    const double thetad = theta * (1.0 + k1 * theta2 + k2 * theta4 + k3 * theta6 + k4 * theta8);
    const double scaling = (r > 1e-8) ? thetad / r : 1.0;
    *d0 = scaling * u0;
    *d1 = scaling * u1;
    if (J) {
      if (r > 1e-8) {
        // "mostly matlab generated" (EquidistantDistortion.hpp:108-150), same temporaries
        double t2, t3, t4, t6, t7, t8, t9, t11, t17, t18, t19, t20, t25;
        t2 = u0 * u0;
        t3 = u1 * u1;
        t4 = t2 + t3;
        t6 = atan(sqrt(t4));
        t7 = t6 * t6;
        t8 = 1.0 / sqrt(t4);
        t9 = t7 * t7;
        t11 = 1.0 / ((t2 + t3) + 1.0);
        t17 = (((k1 * t7 + k2 * t9) + k3 * t7 * t9) + k4 * (t9 * t9)) + 1.0;
        t18 = 1.0 / t4;
        t19 = 1.0 / sqrt(t4 * t4 * t4);
        t20 = t6 * t8 * t17;
        t25 = ((k2 * t6 * t7 * t8 * t11 * u1 * 4.0 + k3 * t6 * t8 * t9 * t11 * u1 * 6.0) +
               k4 * t6 * t7 * t8 * t9 * t11 * u1 * 8.0) +
              k1 * t6 * t8 * t11 * u1 * 2.0;
        t4 = ((k2 * t6 * t7 * t8 * t11 * u0 * 4.0 + k3 * t6 * t8 * t9 * t11 * u0 * 6.0) +
              k4 * t6 * t7 * t8 * t9 * t11 * u0 * 8.0) +
             k1 * t6 * t8 * t11 * u0 * 2.0;
        t7 = t11 * t17 * t18 * u0 * u1;
        J->b = (t7 + t6 * t8 * t25 * u0) - t6 * t17 * t19 * u0 * u1;
        J->d = ((t20 - t3 * t6 * t17 * t19) + t3 * t11 * t17 * t18) + t6 * t8 * t25 * u1;
        J->a = ((t20 - t2 * t6 * t17 * t19) + t2 * t11 * t17 * t18) + t6 * t8 * t4 * u0;
        J->c = (t7 + t6 * t8 * t4 * u1) - t6 * t17 * t19 * u0 * u1;
      } else {
        J->a = 1.0;  // J.setIdentity() (EquidistantDistortion.hpp, r <= 1e-8 branch)
        J->b = 0.0;
        J->c = 0.0;
        J->d = 1.0;
      }
    }
    return true;
  }
  *d0 = u0;  // NoDistortion
  *d1 = u1;
  if (J) {
    J->a = 1.0;
    J->b = 0.0;
    J->c = 0.0;
    J->d = 1.0;
  }
  return true;
}

// D::undistort: Gauss-Newton from the distorted point; success once chi2 < 1e-6 was seen, stop at chi2 < 1e-15.
// radtan: 5 iterations (RadialTangentialDistortion.hpp:200), equidistant: 20 (EquidistantDistortion.hpp:304).
__host__ __device__ inline bool undistort_point(const okc_pinhole& cam, double y0, double y1, double* x0, double* x1) {
  if (cam.distortion == OKC_DISTORTION_NONE) {
    *x0 = y0;
    *x1 = y1;
    return true;
  }
  double xb0 = y0, xb1 = y1;
  const int n = cam.distortion == OKC_DISTORTION_EQUIDISTANT ? 20 : 5;
  bool success = false;
  for (int i = 0; i < n; ++i) {
    double t0, t1;
    Mat2 E;
    distort_point(cam, xb0, xb1, &t0, &t1, &E);
    const double e0 = y0 - t0, e1 = y1 - t1;
    // E2 = E^T * E
    Mat2 E2;
    E2.a = E.a * E.a + E.c * E.c;
    E2.b = E.a * E.b + E.c * E.d;
    E2.c = E.b * E.a + E.d * E.c;
    E2.d = E.b * E.b + E.d * E.d;
    // E2.inverse(): adjugate times 1/det
    const double invdet = 1.0 / (E2.a * E2.d - E2.c * E2.b);
    Mat2 I;
    I.a = E2.d * invdet;
    I.c = -E2.c * invdet;
    I.b = -E2.b * invdet;
    I.d = E2.a * invdet;
    // (E2^-1 * E^T) * e
    Mat2 M;
    M.a = I.a * E.a + I.b * E.b;
    M.b = I.a * E.c + I.b * E.d;
    M.c = I.c * E.a + I.d * E.b;
    M.d = I.c * E.c + I.d * E.d;
    xb0 += M.a * e0 + M.b * e1;
    xb1 += M.c * e0 + M.d * e1;
    const double chi2 = e0 * e0 + e1 * e1;
    if (chi2 < 1e-6) success = true;
    if (chi2 < 1e-15) {
      success = true;
      break;
    }
  }
  *x0 = xb0;
  *x1 = xb1;
  return success;
}

// PinholeCamera::project (PinholeCamera.hpp:249-283). Status codes = cameras::ProjectionStatus (CameraBase.hpp:38-45).
__host__ __device__ inline int project_point(const okc_pinhole& cam, double p0, double p1, double p2, double* u, double* v) {
  if (fabs(p2) < 1.0e-12) return OKC_PROJ_INVALID;
  const double rz = 1.0 / p2;
  double d0, d1;
  if (!distort_point(cam, p0 * rz, p1 * rz, &d0, &d1, nullptr)) return OKC_PROJ_INVALID;
  *u = cam.fu * d0 + cam.cu;
  *v = cam.fv * d1 + cam.cv;
  // isInImage, then isMasked (no mask images here), then the sign of z
  if (*u < 0.0 || *v < 0.0 || *u >= (double)cam.width || *v >= (double)cam.height) return OKC_PROJ_OUTSIDE_IMAGE;
  return p2 > 0.0 ? OKC_PROJ_SUCCESSFUL : OKC_PROJ_BEHIND;
}

__device__ __forceinline__ void back_project_body(const okc_pinhole& cam, const okc_keypoint* __restrict__ kps, int n,
                                                  int k, double* __restrict__ e_C, uint8_t* __restrict__ valid,
                                                  int* __restrict__ nValid) {
  bool ok = false;
  if (k < n) {
    // Eigen::Vector2d(ckp.pt.x, ckp.pt.y): the floats widen exactly
    const double y0 = ((double)kps[k].x - cam.cu) * (1.0 / cam.fu);
    const double y1 = ((double)kps[k].y - cam.cv) * (1.0 / cam.fv);
    double x0, x1;
    ok = undistort_point(cam, y0, y1, &x0, &x1);
    e_C[3 * k] = x0;
    e_C[3 * k + 1] = x1;
    e_C[3 * k + 2] = 1.0;
    valid[k] = ok ? 1 : 0;
  }
  const unsigned m = __ballot_sync(0xffffffffu, ok);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(nValid, __popc(m));
}

__global__ void back_project_kernel(okc_pinhole cam, const okc_keypoint* __restrict__ kps, int n,
                                    double* __restrict__ e_C, uint8_t* __restrict__ valid, int* __restrict__ nValid) {
  back_project_body(cam, kps, n, blockIdx.x * blockDim.x + threadIdx.x, e_C, valid, nValid);
}

// Frame pipeline: Frame::computeBackProjections of every slot's described keypoints, right after describe
// (Frontend::detectAndDescribe, Frontend.cpp:246-248); the count is still on the device.
__global__ void back_project_slots_kernel(const Slot* __restrict__ slots, int maxKp) {
  const Slot& slot = slots[blockIdx.y];
  if (!slot.has_geometry) return;
  const int n = min(slot.counters[2], maxKp);
  if ((int)(blockIdx.x * blockDim.x) >= n) return;
  back_project_body(slot.pinhole, slot.kp_desc, n, blockIdx.x * blockDim.x + threadIdx.x, slot.e_C, slot.bp_valid,
                    slot.counters + 6);
}

// T = {C row-major 3x3, r}: hp_C = [C * hp_W.head<3>() + r * hp_W[3]; hp_W[3]]
__device__ __forceinline__ void project_landmark_body(const okc_pinhole& cam, const double* __restrict__ T_CW,
                                                      const double* __restrict__ hp_W, int l, double reprThreshold,
                                                      double* __restrict__ proj_xy, uint8_t* __restrict__ status,
                                                      uint8_t* __restrict__ visible) {
  const double x = hp_W[4 * l], y = hp_W[4 * l + 1], z = hp_W[4 * l + 2], s = hp_W[4 * l + 3];
  double c[3];
  for (int i = 0; i < 3; ++i)
    c[i] = ((T_CW[3 * i] * x + T_CW[3 * i + 1] * y) + T_CW[3 * i + 2] * z) + T_CW[9 + i] * s;
  // projectHomogeneous: head, negated when the homogeneous coordinate is negative
  if (s < 0) {
    c[0] = -c[0];
    c[1] = -c[1];
    c[2] = -c[2];
  }
  double u = 0.0, v = 0.0;
  const int st = project_point(cam, c[0], c[1], c[2], &u, &v);
  bool vis = !(st == OKC_PROJ_INVALID || st == OKC_PROJ_BEHIND);
  if (vis) {
    const double maxU = (double)cam.width + reprThreshold, maxV = (double)cam.height + reprThreshold;
    if (u < -reprThreshold || v < -reprThreshold || u > maxU || v > maxV) vis = false;
  }
  proj_xy[2 * l] = u;
  proj_xy[2 * l + 1] = v;
  status[l] = (uint8_t)st;
  visible[l] = vis ? 1 : 0;
}

__global__ void project_landmarks_kernel(okc_pinhole cam, const double* __restrict__ T_CW,
                                         const double* __restrict__ hp_W, int n, double reprThreshold,
                                         double* __restrict__ proj_xy, uint8_t* __restrict__ status,
                                         uint8_t* __restrict__ visible) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n) return;
  project_landmark_body(cam, T_CW, hp_W, l, reprThreshold, proj_xy, status, visible);
}

int check_pinhole(const okc_pinhole* cam) {
  if (!cam || cam->width <= 0 || cam->height <= 0) return OKC_ERR_INVALID;
  if (cam->distortion != OKC_DISTORTION_NONE && cam->distortion != OKC_DISTORTION_RADTAN &&
      cam->distortion != OKC_DISTORTION_EQUIDISTANT) {
    set_error("distortion model %d is not implemented (radialtangential8 is used by no shipped configuration)",
              cam->distortion);
    return OKC_ERR_INVALID;
  }
  return OKC_OK;
}

void launch_back_project(const okc_pinhole& cam, const okc_keypoint* kps, int n, double* e_C, uint8_t* valid,
                         int* nValid, cudaStream_t s) {
  if (n > 0) back_project_kernel<<<(n + 127) / 128, 128, 0, s>>>(cam, kps, n, e_C, valid, nValid);
}

void launch_back_project_slots(const Slot* slots, int n, int max_kp, cudaStream_t s) {
  if (n > 0 && max_kp > 0) back_project_slots_kernel<<<dim3((max_kp + 127) / 128, n), 128, 0, s>>>(slots, max_kp);
}

void launch_project_landmarks(const okc_pinhole& cam, const double* T_CW, const double* hp_W, int n,
                              double reprThreshold, double* proj_xy, uint8_t* status, uint8_t* visible,
                              cudaStream_t s) {
  if (n > 0)
    project_landmarks_kernel<<<(n + 127) / 128, 128, 0, s>>>(cam, T_CW, hp_W, n, reprThreshold, proj_xy, status,
                                                              visible);
}


}  // namespace okc

struct okc_context;
int okc_context_device(const okc_context* ctx);

extern "C" {

int okc_back_project(okc_context* ctx, const okc_pinhole* cam, const okc_keypoint* keypoints, int n, double* e_C,
                     uint8_t* valid, int* n_valid) {
  if (!ctx || n < 0 || (n > 0 && (!keypoints || !e_C || !valid))) return OKC_ERR_INVALID;
  if (int rc = okc::check_pinhole(cam)) return rc;
  if (n_valid) *n_valid = 0;
  if (n == 0) return OKC_OK;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  uint8_t* d = nullptr;
  const size_t oKp = 0, oE = (size_t)n * sizeof(okc_keypoint), oV = oE + (size_t)n * 24, oN = (oV + n + 15) & ~(size_t)15;
  OKC_CUDA_TRY(cudaMalloc(&d, oN + 16));
  int rc = OKC_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  };
  fail(cudaMemcpy(d + oKp, keypoints, (size_t)n * sizeof(okc_keypoint), cudaMemcpyHostToDevice), "cudaMemcpy");
  fail(cudaMemset(d + oN, 0, 16), "cudaMemset");
  if (rc == OKC_OK) {
    okc::launch_back_project(*cam, reinterpret_cast<const okc_keypoint*>(d + oKp), n, reinterpret_cast<double*>(d + oE),
                             d + oV, reinterpret_cast<int*>(d + oN), nullptr);
    fail(cudaGetLastError(), "back_project_kernel");
    fail(cudaMemcpy(e_C, d + oE, (size_t)n * 24, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(valid, d + oV, (size_t)n, cudaMemcpyDeviceToHost), "cudaMemcpy");
    int nv = 0;
    fail(cudaMemcpy(&nv, d + oN, sizeof(int), cudaMemcpyDeviceToHost), "cudaMemcpy");
    if (n_valid) *n_valid = nv;
  }
  cudaFree(d);
  return rc;
}

int okc_project_landmarks(okc_context* ctx, const okc_pinhole* cam, const double T_CW[12], const double* hp_W, int n,
                          double repr_threshold, double* proj_xy, uint8_t* status, uint8_t* visible) {
  if (!ctx || !T_CW || n < 0 || (n > 0 && (!hp_W || !proj_xy || !status || !visible))) return OKC_ERR_INVALID;
  if (int rc = okc::check_pinhole(cam)) return rc;
  if (n == 0) return OKC_OK;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  uint8_t* d = nullptr;
  const size_t oT = 0, oH = 128, oP = oH + (size_t)n * 32, oS = oP + (size_t)n * 16, oV = oS + (size_t)n;
  OKC_CUDA_TRY(cudaMalloc(&d, oV + n));
  int rc = OKC_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  };
  fail(cudaMemcpy(d + oT, T_CW, 12 * sizeof(double), cudaMemcpyHostToDevice), "cudaMemcpy");
  fail(cudaMemcpy(d + oH, hp_W, (size_t)n * 32, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (rc == OKC_OK) {
    okc::launch_project_landmarks(*cam, reinterpret_cast<const double*>(d + oT), reinterpret_cast<const double*>(d + oH),
                                  n, repr_threshold, reinterpret_cast<double*>(d + oP), d + oS, d + oV, nullptr);
    fail(cudaGetLastError(), "project_landmarks_kernel");
    fail(cudaMemcpy(proj_xy, d + oP, (size_t)n * 16, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(status, d + oS, (size_t)n, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(visible, d + oV, (size_t)n, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  cudaFree(d);
  return rc;
}

}  // extern "C"

// ----------------------------------------------------------------------------------------------
// Keyframe-overlap rasterisation of Frontend::doWeNeedANewKeyframe (okvis_frontend/src/Frontend.cpp:1186-1295):
// per (frame, camera) two H/10 x W/10 masks - every keypoint / every matched keypoint drawn as a filled circle of
// radius int(min(rows, cols) * 0.09) at cv::Point(keypoint.pt * 0.1) - and the pixel counts of their AND and OR.
// cv::circle(..., cv::FILLED) fills, for each row dy, the columns |dx| <= floor(sqrt(r^2 - dy^2)), clipped to the
// image (pinned against cv2 4.13 goldens, tests/golden/overlap_goldens.npz). One CTA per view, masks in shared memory;
// the stores are idempotent (255), so overlapping circles need no ordering.
// ----------------------------------------------------------------------------------------------
namespace okc {

constexpr int kOverlapThreads = 256;

__global__ void __launch_bounds__(kOverlapThreads)
keyframe_overlap_kernel(const int* __restrict__ width, const int* __restrict__ height,
                        const long long* __restrict__ kpOffset, const okc_keypoint* __restrict__ kps,
                        const uint8_t* __restrict__ matched, int* __restrict__ interCount, int* __restrict__ unionCount) {
  extern __shared__ uint8_t masks[];   // detections | matches
  __shared__ int s_inter, s_union;
  const int view = blockIdx.x;
  const int rows = height[view] / 10, cols = width[view] / 10;
  const int cells = rows * cols;
  uint8_t* det = masks;
  uint8_t* mat = masks + cells;
  for (int i = threadIdx.x; i < 2 * cells; i += kOverlapThreads) masks[i] = 0;
  if (threadIdx.x == 0) s_inter = s_union = 0;
  __syncthreads();
  const int r = (int)((double)(rows < cols ? rows : cols) * 0.09);
  const long long k0 = kpOffset[view], k1 = kpOffset[view + 1];
  for (long long k = k0 + threadIdx.x; k < k1; k += kOverlapThreads) {
    // keypoint.pt * 0.1: Point2f * double -> float; Point2f -> cv::Point rounds half to even (cvRound)
    const int cx = __float2int_rn((float)((double)kps[k].x * 0.1));
    const int cy = __float2int_rn((float)((double)kps[k].y * 0.1));
    const bool m = matched[k] != 0;
    for (int dy = -r; dy <= r; ++dy) {
      const int y = cy + dy;
      if (y < 0 || y >= rows) continue;
      const int hw = (int)sqrt((double)(r * r - dy * dy));
      int xa = cx - hw, xb = cx + hw;
      xa = xa < 0 ? 0 : xa;
      xb = xb >= cols ? cols - 1 : xb;
      for (int x = xa; x <= xb; ++x) {
        det[y * cols + x] = 255;
        if (m) mat[y * cols + x] = 255;
      }
    }
  }
  __syncthreads();
  int ci = 0, cu = 0;
  for (int i = threadIdx.x; i < cells; i += kOverlapThreads) {
    ci += (det[i] & mat[i]) != 0;
    cu += (det[i] | mat[i]) != 0;
  }
  for (int d = 16; d > 0; d >>= 1) {
    ci += __shfl_down_sync(0xffffffffu, ci, d);
    cu += __shfl_down_sync(0xffffffffu, cu, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&s_inter, ci);
    atomicAdd(&s_union, cu);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    interCount[view] = s_inter;
    unionCount[view] = s_union;
  }
}

}  // namespace okc

extern "C" int okc_keyframe_overlap(okc_context* ctx, int n_views, const int* width, const int* height,
                                    const int64_t* kp_offset, const okc_keypoint* keypoints, const uint8_t* matched,
                                    int32_t* intersection, int32_t* uni) {
  if (!ctx || n_views < 0 || (n_views > 0 && (!width || !height || !kp_offset || !intersection || !uni)))
    return OKC_ERR_INVALID;
  if (n_views == 0) return OKC_OK;
  const int64_t nk = kp_offset[n_views];
  if (kp_offset[0] != 0 || nk < 0 || (nk > 0 && (!keypoints || !matched))) return OKC_ERR_INVALID;
  size_t smem = 0;
  for (int v = 0; v < n_views; ++v) {
    if (width[v] < 10 || height[v] < 10 || kp_offset[v + 1] < kp_offset[v]) return OKC_ERR_INVALID;
    const size_t need = 2 * (size_t)(width[v] / 10) * (size_t)(height[v] / 10);
    smem = need > smem ? need : smem;
  }
  if (smem > 200 * 1024) {
    okc::set_error("okc_keyframe_overlap: image too large for the shared-memory masks");
    return OKC_ERR_INVALID;
  }
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  const size_t oW = 0, oH = oW + (size_t)n_views * 4, oO = (oH + (size_t)n_views * 4 + 7) & ~(size_t)7,
               oK = oO + (size_t)(n_views + 1) * 8, oM = oK + (size_t)nk * sizeof(okc_keypoint),
               oI = (oM + (size_t)nk + 7) & ~(size_t)7, oU = oI + (size_t)n_views * 4, total = oU + (size_t)n_views * 4;
  uint8_t* d = nullptr;
  OKC_CUDA_TRY(cudaMalloc(&d, total));
  int rc = OKC_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  };
  fail(cudaMemcpy(d + oW, width, (size_t)n_views * 4, cudaMemcpyHostToDevice), "cudaMemcpy");
  fail(cudaMemcpy(d + oH, height, (size_t)n_views * 4, cudaMemcpyHostToDevice), "cudaMemcpy");
  fail(cudaMemcpy(d + oO, kp_offset, (size_t)(n_views + 1) * 8, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (nk > 0) {
    fail(cudaMemcpy(d + oK, keypoints, (size_t)nk * sizeof(okc_keypoint), cudaMemcpyHostToDevice), "cudaMemcpy");
    fail(cudaMemcpy(d + oM, matched, (size_t)nk, cudaMemcpyHostToDevice), "cudaMemcpy");
  }
  if (rc == OKC_OK) {
    if (smem > 48 * 1024)
      fail(cudaFuncSetAttribute(okc::keyframe_overlap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
           "cudaFuncSetAttribute");
    okc::keyframe_overlap_kernel<<<n_views, okc::kOverlapThreads, smem>>>(
        reinterpret_cast<const int*>(d + oW), reinterpret_cast<const int*>(d + oH),
        reinterpret_cast<const long long*>(d + oO), reinterpret_cast<const okc_keypoint*>(d + oK), d + oM,
        reinterpret_cast<int*>(d + oI), reinterpret_cast<int*>(d + oU));
    fail(cudaGetLastError(), "keyframe_overlap_kernel");
    fail(cudaMemcpy(intersection, d + oI, (size_t)n_views * 4, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(uni, d + oU, (size_t)n_views * 4, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  cudaFree(d);
  return rc;
}

// ----------------------------------------------------------------------------------------------
// Landmark preparation of Frontend::matchToMap, second half (okvis_frontend/src/Frontend.cpp:1409-1500, SURVEY 8(f)
// row f1): per visible landmark walk its observations (newest first), decide is3d (:1433-1439), drop observations
// with more than 0.6 rad viewpoint change or 50 % scale change (:1442-1453), keep the best 3 by
// score = 0.5 * (acos(cosViewpointChange) / 0.6 + scaleChange / 0.5) with the reference's "replace the worst slot"
// rule (:1455-1487), and pack the kept descriptor rows, world rays and camera centres into the pool that
// matchToMapByThread / matchToMapByThreadUnitialised sweep (rows of a landmark contiguous, landmarks ascending).
// One thread per landmark; Eigen's 3-vector expressions written out ((x0 y0 + x1 y1) + x2 y2, normalized() = v /
// sqrt(v.v) when v.v > 0). acos / cos come from the device math library: scores can differ from glibc's in the last
// ulp, which matters only for exact ties between observation scores.
// ----------------------------------------------------------------------------------------------
namespace okc {

__device__ __forceinline__ double dot3(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
__device__ __forceinline__ void normalized3(const double v[3], double o[3]) {
  const double n2 = dot3(v, v);
  if (n2 > 0.0) {
    const double n = sqrt(n2);
    o[0] = v[0] / n;
    o[1] = v[1] / n;
    o[2] = v[2] / n;
  } else {
    o[0] = v[0];
    o[1] = v[1];
    o[2] = v[2];
  }
}

__device__ __forceinline__ void prune_observations_body(const PrepareArgs& a, int l, const double* __restrict__ r_WC1) {
  int* so = a.slot_obs + 3 * l;
  so[0] = so[1] = so[2] = -1;
  a.is3d[l] = 0;
  a.rows[l] = 0;
  if (!a.visible[l]) return;
  const double* hp = a.hp_W + 4 * l;
  const double p_W[3] = {hp[0] / hp[3], hp[1] / hp[3], hp[2] / hp[3]};
  const double r_W[3] = {p_W[0] - r_WC1[0], p_W[1] - r_WC1[1], p_W[2] - r_WC1[2]};
  double e_W[3];
  normalized3(r_W, e_W);
  const double rn = sqrt(dot3(r_W, r_W));
  const double r = rn > 0.01 ? rn : 0.01;   // std::max(0.01, r_W.norm())
  const double quality = a.quality[l];
  const double cos06 = cos(0.6), cos3d = cos(10.0 / a.focal_sum);
  double best[3] = {1.0, 1.0, 1.0};
  bool is3d = false;
  int o = 0;
  for (long long i = a.obs_offset[l]; i < a.obs_offset[l + 1]; ++i) {
    const int v = a.obs_view[i];
    const double* ro = a.view_r + 3 * v;
    const double r_W_old[3] = {p_W[0] - ro[0], p_W[1] - ro[1], p_W[2] - ro[2]};
    if (!is3d) {
      const double s = 0.2 / a.focal_sum / quality;
      const double r_close[3] = {r_W[0] - s * r_W_old[0], r_W[1] - s * r_W_old[1], r_W[2] - s * r_W_old[2]};
      double n1[3], n2[3];
      normalized3(r_W, n1);
      normalized3(r_close, n2);
      if (dot3(n1, n2) > cos3d) is3d = true;
    }
    double eo[3];
    normalized3(r_W_old, eo);
    const double cosV = dot3(e_W, eo);
    if (cosV < cos06 && !a.loop_closure) continue;
    const double scaleChange = fabs(r - sqrt(dot3(r_W_old, r_W_old))) / r;
    if (scaleChange > 0.5 && !a.loop_closure) continue;
    const double score = 0.5 * (acos(cosV) / 0.6 + scaleChange / 0.5);
    double worst = 0.0;
    int worstIdx = 0;
    for (int n = 0; n < 3; ++n) {
      if (best[n] > worst) {
        worst = best[n];
        worstIdx = n;
      }
    }
    if (score < best[worstIdx]) {
      so[worstIdx] = (int)(i - a.obs_offset[l]);
      o = o > worstIdx ? o : worstIdx;
      best[worstIdx] = score;
    }
  }
  a.is3d[l] = is3d ? 1 : 0;
  a.rows[l] = o + 1;
}

__global__ void prune_observations_kernel(PrepareArgs a) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l < a.L) prune_observations_body(a, l, a.r_WC1);
}

// exclusive scan of rows[] -> row_offset[L + 1] (one block; L is a few thousand)
__device__ __forceinline__ void scan_rows_body(const int* __restrict__ rows, int L, int* __restrict__ offset) {
  __shared__ int s_part[1024];
  const int tid = threadIdx.x;
  const int per = (L + 1023) / 1024;
  const int b = tid * per, e = min(L, b + per);
  int sum = 0;
  for (int i = b; i < e; ++i) sum += rows[i];
  s_part[tid] = sum;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const int t = tid >= d ? s_part[tid - d] : 0;
    __syncthreads();
    s_part[tid] += t;
    __syncthreads();
  }
  int run = s_part[tid] - sum;
  for (int i = b; i < e; ++i) {
    offset[i] = run;
    run += rows[i];
  }
  if (tid == 1023) offset[L] = s_part[1023];
}

__global__ void __launch_bounds__(1024) scan_rows_kernel(const int* __restrict__ rows, int L, int* __restrict__ offset) {
  scan_rows_body(rows, L, offset);
}

// pool rows: descriptor, owner (landmark index), e_W = C_old * e_C.normalized(), r_W = r_old, filled flag
__device__ __forceinline__ void pack_pool_body(const PrepareArgs& a, int t, const int* __restrict__ offset,
                                               uint8_t* __restrict__ pool, int* __restrict__ owner,
                                               double* __restrict__ e_W, double* __restrict__ r_W,
                                               uint8_t* __restrict__ filled, int* __restrict__ row_obs) {
  const int l = t / 3, n = t - 3 * l;
  if (l >= a.L || n >= a.rows[l]) return;
  const int row = offset[l] + n;
  const int so = a.slot_obs[3 * l + n];
  owner[row] = l;
  filled[row] = so >= 0;
  row_obs[row] = so;
  uint4* dst = reinterpret_cast<uint4*>(pool) + (size_t)row * 3;
  if (so < 0) {   // the reference leaves such a row uninitialised (cv::Mat / Eigen resize); zeros here
    dst[0] = dst[1] = dst[2] = make_uint4(0, 0, 0, 0);
    for (int i = 0; i < 3; ++i) e_W[3 * row + i] = r_W[3 * row + i] = 0.0;
    return;
  }
  const long long i = a.obs_offset[l] + so;
  const int v = a.obs_view[i];
  const long long src = a.view_row0[v] + a.obs_kp[i];
  const uint4* s4 = reinterpret_cast<const uint4*>(a.desc) + (size_t)src * 3;
  dst[0] = s4[0];
  dst[1] = s4[1];
  dst[2] = s4[2];
  const double ec[3] = {a.e_C[3 * src], a.e_C[3 * src + 1], a.e_C[3 * src + 2]};
  double en[3];
  normalized3(ec, en);
  const double* C = a.view_C + 9 * v;
  for (int r = 0; r < 3; ++r) {
    e_W[3 * row + r] = (C[3 * r] * en[0] + C[3 * r + 1] * en[1]) + C[3 * r + 2] * en[2];
    r_W[3 * row + r] = a.view_r[3 * v + r];
  }
}

__global__ void pack_pool_kernel(PrepareArgs a, const int* __restrict__ offset, uint8_t* __restrict__ pool,
                                 int* __restrict__ owner, double* __restrict__ e_W, double* __restrict__ r_W,
                                 uint8_t* __restrict__ filled, int* __restrict__ row_obs) {
  pack_pool_body(a, blockIdx.x * blockDim.x + threadIdx.x, offset, pool, owner, e_W, r_W, filled, row_obs);
}

// Frame pipeline (okc_batch_prepare_landmarks_enqueue): the same three steps plus the projection for every slot of a
// batch at once, blockIdx.y = slot. The landmark pool (points, observations, descriptors of the keyframe window) stays
// in HBM; per frame only the slot's pose changes.
__global__ void project_landmarks_slots_kernel(const PrepSlot* __restrict__ prep) {
  const PrepSlot& p = prep[blockIdx.y];
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l < p.a.L)
    project_landmark_body(p.cam, p.pose_slot ? p.pose_slot->T_CW : p.T_CW, p.a.hp_W, l, p.repr_threshold, p.proj_xy,
                          p.status, p.visible);
}
__global__ void prune_observations_slots_kernel(const PrepSlot* __restrict__ prep) {
  const PrepSlot& p = prep[blockIdx.y];
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l < p.a.L) prune_observations_body(p.a, l, p.pose_slot ? p.pose_slot->r_WC : p.a.r_WC1);
}
__global__ void __launch_bounds__(1024) scan_rows_slots_kernel(const PrepSlot* __restrict__ prep) {
  const PrepSlot& p = prep[blockIdx.x];
  scan_rows_body(p.a.rows, p.a.L, p.row_offset);
  // the rows of the landmarks that are not 3D (what matchToMapByThreadUnitialised sweeps), counted the same way:
  // u_count[l] = is3d[l] ? 0 : rows[l] is written over slot_obs' neighbour scratch first
  __syncthreads();
  for (int l = threadIdx.x; l < p.a.L; l += 1024) p.u_count[l] = p.a.is3d[l] ? 0 : p.a.rows[l];
  __syncthreads();
  scan_rows_body(p.u_count, p.a.L, p.u_offset);
}
__global__ void pack_pool_slots_kernel(const PrepSlot* __restrict__ prep) {
  const PrepSlot& p = prep[blockIdx.y];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  pack_pool_body(p.a, t, p.row_offset, p.pool, p.owner, p.e_W, p.r_W, p.filled, p.row_obs);
  const int l = t / 3, n = t - 3 * l;
  if (l < p.a.L && n < p.a.rows[l] && !p.a.is3d[l]) {
    const int row = p.row_offset[l] + n, pos = p.u_offset[l] + n;
    p.u_rows[pos] = row;
    // (this thread wrote pool row `row` just above)
    const uint4* src = reinterpret_cast<const uint4*>(p.pool) + (size_t)row * 3;
    uint4* dst = reinterpret_cast<uint4*>(p.u_desc) + (size_t)pos * 3;
    dst[0] = src[0];
    dst[1] = src[1];
    dst[2] = src[2];
  }
}

void launch_prepare_landmarks_slots(const PrepSlot* prep, int n, int L, cudaStream_t s) {
  if (n <= 0 || L <= 0) return;
  project_landmarks_slots_kernel<<<dim3((L + 127) / 128, n), 128, 0, s>>>(prep);
  prune_observations_slots_kernel<<<dim3((L + 127) / 128, n), 128, 0, s>>>(prep);
  scan_rows_slots_kernel<<<n, 1024, 0, s>>>(prep);
  pack_pool_slots_kernel<<<dim3((3 * L + 127) / 128, n), 128, 0, s>>>(prep);
}

}  // namespace okc

extern "C" int okc_prepare_landmarks(okc_context* ctx, const okc_landmark_views* views, int n_landmarks,
                                     const double* hp_W, const double* quality, const uint8_t* visible,
                                     const int64_t* obs_offset, const int32_t* obs_view, const int32_t* obs_kp,
                                     const double r_WC1[3], double focal_sum, int loop_closure, uint8_t* is3d,
                                     int32_t* rows, int32_t* nm, uint8_t* pool, int32_t* owner, double* e_W,
                                     double* r_W, uint8_t* filled, int32_t* row_obs) {
  const int L = n_landmarks;
  if (!ctx || !views || L < 0 || !r_WC1 || !nm) return OKC_ERR_INVALID;
  *nm = 0;
  if (L == 0) return OKC_OK;
  if (!hp_W || !quality || !visible || !obs_offset || !is3d || !rows || !pool || !owner || !e_W || !r_W || !filled ||
      !row_obs || views->n_views < 0 || views->n_rows < 0)
    return OKC_ERR_INVALID;
  const int64_t nObs = obs_offset[L];
  if (obs_offset[0] != 0 || nObs < 0 || (nObs > 0 && (!obs_view || !obs_kp))) return OKC_ERR_INVALID;
  for (int l = 0; l < L; ++l)
    if (obs_offset[l + 1] < obs_offset[l]) return OKC_ERR_INVALID;
  const int V = views->n_views;
  for (int64_t i = 0; i < nObs; ++i) {
    const int v = obs_view[i];
    if (v < 0 || v >= V) return OKC_ERR_INVALID;
    const int64_t end = v + 1 < V ? views->row0[v + 1] : views->n_rows;
    if (obs_kp[i] < 0 || views->row0[v] + obs_kp[i] >= end) {
      okc::set_error("okc_prepare_landmarks: observation %lld points outside its view", (long long)i);
      return OKC_ERR_INVALID;
    }
  }
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  // one device block: inputs, then outputs
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t o = off;
    off = (off + bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t N = (size_t)views->n_rows;
  const size_t oHp = take((size_t)L * 32), oQ = take((size_t)L * 8), oVis = take((size_t)L), oOff = take((size_t)(L + 1) * 8),
               oOv = take((size_t)nObs * 4), oOk = take((size_t)nObs * 4), oC = take((size_t)V * 72),
               oR = take((size_t)V * 24), oRow0 = take((size_t)V * 8), oDesc = take(N * 48), oEc = take(N * 24),
               oIs3d = take((size_t)L), oRows = take((size_t)L * 4), oSlot = take((size_t)L * 12),
               oRowOff = take((size_t)(L + 1) * 4), oPool = take((size_t)L * 3 * 48), oOwner = take((size_t)L * 12),
               oEw = take((size_t)L * 72), oRw = take((size_t)L * 72), oFill = take((size_t)L * 3),
               oRobs = take((size_t)L * 12);
  uint8_t* d = nullptr;
  OKC_CUDA_TRY(cudaMalloc(&d, off));
  int rc = OKC_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  };
  auto up = [&](size_t o, const void* src, size_t bytes) {
    if (bytes) fail(cudaMemcpy(d + o, src, bytes, cudaMemcpyHostToDevice), "cudaMemcpy");
  };
  up(oHp, hp_W, (size_t)L * 32);
  up(oQ, quality, (size_t)L * 8);
  up(oVis, visible, (size_t)L);
  up(oOff, obs_offset, (size_t)(L + 1) * 8);
  up(oOv, obs_view, (size_t)nObs * 4);
  up(oOk, obs_kp, (size_t)nObs * 4);
  up(oC, views->C_WC, (size_t)V * 72);
  up(oR, views->r_WC, (size_t)V * 24);
  up(oRow0, views->row0, (size_t)V * 8);
  up(oDesc, views->desc, N * 48);
  up(oEc, views->e_C, N * 24);
  if (rc == OKC_OK) {
    okc::PrepareArgs a;
    a.hp_W = reinterpret_cast<const double*>(d + oHp);
    a.quality = reinterpret_cast<const double*>(d + oQ);
    a.visible = d + oVis;
    a.obs_offset = reinterpret_cast<const long long*>(d + oOff);
    a.obs_view = reinterpret_cast<const int*>(d + oOv);
    a.obs_kp = reinterpret_cast<const int*>(d + oOk);
    a.view_C = reinterpret_cast<const double*>(d + oC);
    a.view_r = reinterpret_cast<const double*>(d + oR);
    a.view_row0 = reinterpret_cast<const long long*>(d + oRow0);
    a.desc = d + oDesc;
    a.e_C = reinterpret_cast<const double*>(d + oEc);
    for (int i = 0; i < 3; ++i) a.r_WC1[i] = r_WC1[i];
    a.focal_sum = focal_sum;
    a.loop_closure = loop_closure ? 1 : 0;
    a.L = L;
    a.is3d = d + oIs3d;
    a.rows = reinterpret_cast<int*>(d + oRows);
    a.slot_obs = reinterpret_cast<int*>(d + oSlot);
    okc::prune_observations_kernel<<<(L + 127) / 128, 128>>>(a);
    okc::scan_rows_kernel<<<1, 1024>>>(a.rows, L, reinterpret_cast<int*>(d + oRowOff));
    okc::pack_pool_kernel<<<(3 * L + 127) / 128, 128>>>(a, reinterpret_cast<const int*>(d + oRowOff), d + oPool,
                                                        reinterpret_cast<int*>(d + oOwner),
                                                        reinterpret_cast<double*>(d + oEw),
                                                        reinterpret_cast<double*>(d + oRw), d + oFill,
                                                        reinterpret_cast<int*>(d + oRobs));
    fail(cudaGetLastError(), "okc_prepare_landmarks kernels");
    int total = 0;
    fail(cudaMemcpy(&total, d + oRowOff + (size_t)L * 4, 4, cudaMemcpyDeviceToHost), "cudaMemcpy");
    auto down = [&](void* dst, size_t o, size_t bytes) {
      if (bytes) fail(cudaMemcpy(dst, d + o, bytes, cudaMemcpyDeviceToHost), "cudaMemcpy");
    };
    down(is3d, oIs3d, (size_t)L);
    down(rows, oRows, (size_t)L * 4);
    if (rc == OKC_OK) {
      down(pool, oPool, (size_t)total * 48);
      down(owner, oOwner, (size_t)total * 4);
      down(e_W, oEw, (size_t)total * 24);
      down(r_W, oRw, (size_t)total * 24);
      down(filled, oFill, (size_t)total);
      down(row_obs, oRobs, (size_t)total * 4);
      *nm = total;
    }
  }
  cudaFree(d);
  return rc;
}

// ----------------------------------------------------------------------------------------------
// Place-recognition front half: DBoW2::TemplatedVocabulary<FBrisk>::transform of one feature
// (external/DBoW2/include/DBoW2/TemplatedVocabulary.h:1216-1257) - descend the vocabulary tree, at every level the
// child with the smallest FBrisk::distance (= brisk::Hamming::PopcntofXORed over 48 bytes,
// okvis_frontend/src/FBrisk.cpp:62-65), first child wins ties (strict '<') - for all descriptors of a frame at once.
// Frontend::getFilteredDBoWResult / database.add (Frontend.cpp:605-671, 791-806) call it through
// TemplatedDatabase::query / add (TemplatedDatabase.h:409-420, 571). One thread per feature; the tree (819 nodes of
// 48 bytes for resources/small_voc.yml.gz) stays in L1/L2.
// ----------------------------------------------------------------------------------------------
struct okc_vocabulary {
  okc_context* ctx = nullptr;
  int n_nodes = 0;            // including the root (node id 0)
  int* child_offset = nullptr;   // n_nodes + 1
  int* children = nullptr;       // node ids, per parent in file order
  uint4* desc = nullptr;         // n_nodes x 3
  double* weight = nullptr;      // n_nodes
  int* word = nullptr;           // n_nodes: word id of a leaf, -1 otherwise
};

namespace okc {

__global__ void vocabulary_transform_kernel(const int* __restrict__ childOffset, const int* __restrict__ children,
                                            const uint4* __restrict__ nodeDesc, const double* __restrict__ nodeWeight,
                                            const int* __restrict__ nodeWord, const uint4* __restrict__ features, int n,
                                            int nidLevel, unsigned* __restrict__ wordId, double* __restrict__ weight,
                                            unsigned* __restrict__ nodeId) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n) return;
  const uint4 q0 = features[3 * f], q1 = features[3 * f + 1], q2 = features[3 * f + 2];
  int final_id = 0, current_level = 0;
  unsigned nid = 0;   // "if (nid_level <= 0) *nid = 0" and the default when the level is never reached
  do {
    ++current_level;
    const int b = childOffset[final_id], e = childOffset[final_id + 1];
    unsigned best = 0xffffffffu;
    int bestId = children[b];
    for (int c = b; c < e; ++c) {
      const int id = children[c];
      const uint4 d0 = __ldg(nodeDesc + 3 * id), d1 = __ldg(nodeDesc + 3 * id + 1), d2 = __ldg(nodeDesc + 3 * id + 2);
      const unsigned d = __popc(q0.x ^ d0.x) + __popc(q0.y ^ d0.y) + __popc(q0.z ^ d0.z) + __popc(q0.w ^ d0.w) +
                         __popc(q1.x ^ d1.x) + __popc(q1.y ^ d1.y) + __popc(q1.z ^ d1.z) + __popc(q1.w ^ d1.w) +
                         __popc(q2.x ^ d2.x) + __popc(q2.y ^ d2.y) + __popc(q2.z ^ d2.z) + __popc(q2.w ^ d2.w);
      if (d < best) {   // the first child initialises, later ones need a strictly smaller distance
        best = d;
        bestId = id;
      }
    }
    final_id = bestId;
    if (current_level == nidLevel) nid = (unsigned)final_id;
  } while (childOffset[final_id + 1] > childOffset[final_id]);   // !isLeaf()
  wordId[f] = (unsigned)nodeWord[final_id];
  weight[f] = nodeWeight[final_id];
  nodeId[f] = nid;
}

}  // namespace okc

extern "C" {

int okc_vocabulary_destroy(okc_vocabulary* v) {
  if (!v) return OKC_OK;
  cudaFree(v->child_offset);
  cudaFree(v->children);
  cudaFree(v->desc);
  cudaFree(v->weight);
  cudaFree(v->word);
  delete v;
  return OKC_OK;
}

int okc_vocabulary_create(okc_context* ctx, int n, const int32_t* node_id, const int32_t* parent_id,
                          const uint8_t* descriptors, const double* weights, const int32_t* word_of_node, int depth,
                          okc_vocabulary** out) {
  if (!ctx || !out || n <= 0 || !node_id || !parent_id || !descriptors || !weights || !word_of_node || depth <= 0)
    return OKC_ERR_INVALID;
  *out = nullptr;
  const int N = n + 1;   // + root
  std::vector<int> count(N, 0), offset(N + 1, 0), children(n), cursor, word(N, -1);
  std::vector<uint8_t> desc((size_t)N * 48, 0);
  std::vector<double> w(N, 0.0);
  std::vector<char> seen(N, 0);
  for (int i = 0; i < n; ++i) {
    const int id = node_id[i], p = parent_id[i];
    if (id <= 0 || id >= N || p < 0 || p >= N || seen[id]) {
      okc::set_error("okc_vocabulary_create: node ids must be a permutation of 1..n, parents in 0..n");
      return OKC_ERR_INVALID;
    }
    seen[id] = 1;
    ++count[p];
  }
  for (int p = 0; p < N; ++p) offset[p + 1] = offset[p] + count[p];
  cursor.assign(offset.begin(), offset.end() - 1);
  for (int i = 0; i < n; ++i) {   // file order = the order children were appended on load (TemplatedVocabulary.h load())
    const int id = node_id[i];
    children[cursor[parent_id[i]]++] = id;
    memcpy(&desc[(size_t)id * 48], descriptors + (size_t)i * 48, 48);
    w[id] = weights[i];
    word[id] = word_of_node[i];
  }
  if (count[0] == 0) return OKC_ERR_INVALID;
  // every root-to-leaf walk must end within `depth` levels at a node with a word id (guards the device loop)
  std::vector<int> parentOf(N, 0);
  for (int i = 0; i < n; ++i) parentOf[node_id[i]] = parent_id[i];
  for (int id = 1; id < N; ++id) {
    int d = 0, x = id;
    while (x != 0 && d <= depth) {
      x = parentOf[x];
      ++d;
    }
    if (x != 0 || (count[id] == 0 && word[id] < 0)) {
      okc::set_error("okc_vocabulary_create: node %d is deeper than %d levels, detached, or a leaf without a word", id, depth);
      return OKC_ERR_INVALID;
    }
  }
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(ctx)));
  okc_vocabulary* v = new (std::nothrow) okc_vocabulary();
  if (!v) return OKC_ERR_CUDA;
  v->ctx = ctx;
  v->n_nodes = N;
  cudaError_t e = cudaSuccess;
  auto put = [&](void** dst, const void* src, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(dst, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
  };
  put(reinterpret_cast<void**>(&v->child_offset), offset.data(), (size_t)(N + 1) * 4);
  put(reinterpret_cast<void**>(&v->children), children.data(), (size_t)n * 4);
  put(reinterpret_cast<void**>(&v->desc), desc.data(), (size_t)N * 48);
  put(reinterpret_cast<void**>(&v->weight), w.data(), (size_t)N * 8);
  put(reinterpret_cast<void**>(&v->word), word.data(), (size_t)N * 4);
  if (e != cudaSuccess) {
    okc::set_error("okc_vocabulary_create: %s", cudaGetErrorString(e));
    okc_vocabulary_destroy(v);
    return OKC_ERR_CUDA;
  }
  *out = v;
  return OKC_OK;
}

int okc_vocabulary_transform(okc_vocabulary* v, const uint8_t* features, int n, int levels, int levelsup,
                             uint32_t* word_id, double* weight, uint32_t* node_id) {
  if (!v || n < 0 || (n > 0 && (!features || !word_id || !weight || !node_id))) return OKC_ERR_INVALID;
  if (n == 0) return OKC_OK;
  OKC_CUDA_TRY(cudaSetDevice(okc_context_device(v->ctx)));
  uint8_t* d = nullptr;
  const size_t oF = 0, oW = (size_t)n * 48, oWt = oW + (((size_t)n * 4 + 7) & ~(size_t)7), oN = oWt + (size_t)n * 8;
  OKC_CUDA_TRY(cudaMalloc(&d, oN + (size_t)n * 4));
  int rc = OKC_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc == OKC_OK) {
      okc::set_error("%s failed: %s", what, cudaGetErrorString(e));
      rc = OKC_ERR_CUDA;
    }
  };
  fail(cudaMemcpy(d + oF, features, (size_t)n * 48, cudaMemcpyHostToDevice), "cudaMemcpy");
  if (rc == OKC_OK) {
    okc::vocabulary_transform_kernel<<<(n + 127) / 128, 128>>>(
        v->child_offset, v->children, v->desc, v->weight, v->word, reinterpret_cast<const uint4*>(d + oF), n,
        levels - levelsup, reinterpret_cast<unsigned*>(d + oW), reinterpret_cast<double*>(d + oWt),
        reinterpret_cast<unsigned*>(d + oN));
    fail(cudaGetLastError(), "vocabulary_transform_kernel");
    fail(cudaMemcpy(word_id, d + oW, (size_t)n * 4, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(weight, d + oWt, (size_t)n * 8, cudaMemcpyDeviceToHost), "cudaMemcpy");
    fail(cudaMemcpy(node_id, d + oN, (size_t)n * 4, cudaMemcpyDeviceToHost), "cudaMemcpy");
  }
  cudaFree(d);
  return rc;
}

}  // extern "C"
