// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
//
// Oracle for SURVEY.md §8(f) row N2 (per-frame glue right AFTER ORBextractor::operator()):
//   Frame::UndistortKeyPoints      src/Frame.cc:886-913  -> cv::undistortPoints(mat, mat, K, D, cv::Mat(), K)
//   Frame::ComputeStereoFromRGBD   src/Frame.cc:1115-1133
// cv::undistortPoints (calib3d undistort.dispatch.cpp, cvUndistortPointsInternal) restated: float
// inputs promoted to double, 5 fixed-point iterations (default TermCriteria(MAX_ITER, 5, 0.01)) of
//   r2 = x*x + y*y;  icdist = (1 + ((k7*r2 + k6)*r2 + k5)*r2) / (1 + ((k4*r2 + k1)*r2 + k0)*r2)
//   dX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2;  dY = k2*(r2 + 2*y*y) + 2*k3*x*y + k10*r2 + k11*r2*r2
//   x = (x0 - dX)*icdist;  y = (y0 - dY)*icdist
// then the projection by P = K, rounded to float.  Pinned against cv2.undistortPoints 4.13.0
// (tests/test_oracle_frameglue.py, tests/golden/frameglue_cv2.npz).  Compiled -ffp-contract=off.
#include <cmath>
#include <cstring>

#include "../include/cuorb.h"  // POD layout of cuorb_frustum_camera only
#include "orb_oracle.hpp"

namespace orb_oracle {

// K = (fx, fy, cx, cy) as the float entries of mK; dist = k1, k2, p1, p2, k3 (float, k3 = 0 if absent)
void undistort_keypoints(const KeyPoint* in, KeyPoint* out, int n, const float K[4], const float dist[5]) {
  if (dist[0] == 0.0f) {  // src/Frame.cc:887-890
    if (out != in) memcpy(out, in, sizeof(KeyPoint) * (size_t)n);
    return;
  }
  const double fx = K[0], fy = K[1], cx = K[2], cy = K[3], ifx = 1. / fx, ify = 1. / fy;
  double k[14] = {0};
  for (int i = 0; i < 5; i++) k[i] = dist[i];
  for (int i = 0; i < n; i++) {
    double x = in[i].x, y = in[i].y;
    const double u = x, v = y;
    x = (x - cx) * ifx; y = (y - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    // RR = P(3x3) * I with P = K
    const double xx = fx * x + 0. * y + cx, yy = 0. * x + fy * y + cy, ww = 1. / (0. * x + 0. * y + 1.);
    out[i] = in[i];
    out[i].x = (float)(xx * ww);
    out[i].y = (float)(yy * ww);
  }
}

// depth: CV_32F image (imDepth.at<float>(v, u) with the float coordinates truncated to int)
void stereo_from_rgbd(const KeyPoint* kps, const KeyPoint* kps_un, int n, const float* depth, size_t stride_floats,
                      float mbf, float* u_right, float* depth_out) {
  for (int i = 0; i < n; i++) {
    u_right[i] = -1; depth_out[i] = -1;
    const float d = depth[(size_t)(int)kps[i].y * stride_floats + (int)kps[i].x];
    if (d > 0) {
      depth_out[i] = d;
      u_right[i]   = kps_un[i].x - mbf / d;
    }
  }
}

// Frame::isInFrustum, Nleft == -1 branch (src/Frame.cc:644-707) with MapPoint::PredictScale
// (src/MapPoint.cc:574-589) and Pinhole::project (src/CameraModels/Pinhole.cpp:70-76), for n points.
// Float expressions as written, one rounding per operation (-ffp-contract=off); sums of three products
// associate as Eigen 3.4's unrolled size-3 reduction does, x0 + (x1 + x2).  Checked against the reference's
// own lifted body in tests/test_oracle_vs_ref.py.
int is_in_frustum(const cuorb_frustum_camera& c, int n, const float* P3, const float* N3, const float* minD,
                  const float* maxD, float viewingCosLimit, uint8_t* in_view, float* projX, float* projY,
                  float* projXR, int32_t* level, float* viewCosOut, float* depth) {
  int count = 0;
  for (int i = 0; i < n; i++) {
    in_view[i] = 0; projX[i] = -1; projY[i] = -1;              // :646-648
    projXR[i] = -1; level[i] = -1; viewCosOut[i] = 0; depth[i] = 0;
    const float* P = P3 + 3 * i;
    float Pc[3];
    for (int r = 0; r < 3; r++)                                  // Pc = mRcw * P + mtcw (:654)
      Pc[r] = (c.rcw[3 * r] * P[0] + (c.rcw[3 * r + 1] * P[1] + c.rcw[3 * r + 2] * P[2])) + c.tcw[r];
    const float Pc_dist = std::sqrt(Pc[0] * Pc[0] + (Pc[1] * Pc[1] + Pc[2] * Pc[2]));   // :655
    const float PcZ = Pc[2];
    const float invz = 1.0f / PcZ;                               // :659
    if (PcZ < 0.0f) continue;                                    // :660
    const float u = c.fx * Pc[0] / Pc[2] + c.cx, v = c.fy * Pc[1] / Pc[2] + c.cy;   // Pinhole::project
    if (u < c.min_x || u > c.max_x) continue;                    // :666
    if (v < c.min_y || v > c.max_y) continue;                    // :669
    projX[i] = u; projY[i] = v;                                  // :673-674
    const float maxDistance = 1.2f * maxD[i], minDistance = 0.8f * minD[i];   // MapPoint.cc:547-555
    const float PO[3] = {P[0] - c.ow[0], P[1] - c.ow[1], P[2] - c.ow[2]};
    const float dist = std::sqrt(PO[0] * PO[0] + (PO[1] * PO[1] + PO[2] * PO[2]));     // :680
    if (dist < minDistance || dist > maxDistance) continue;      // :682
    const float* Pn = N3 + 3 * i;
    const float viewCos = (PO[0] * Pn[0] + (PO[1] * Pn[1] + PO[2] * Pn[2])) / dist;    // :689
    if (viewCos < viewingCosLimit) continue;                     // :691
    const float ratio = maxD[i] / dist;                          // MapPoint.cc:578
    int nScale = (int)std::ceil(std::log(ratio) / c.log_scale_factor);   // logf, float division (:581)
    if (nScale < 0) nScale = 0;
    else if (nScale >= c.n_levels) nScale = c.n_levels - 1;
    in_view[i] = 1; count++;
    projXR[i] = u - c.mbf * invz;                                // :701
    depth[i] = Pc_dist; level[i] = nScale; viewCosOut[i] = viewCos;
  }
  return count;
}

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {
void oo_undistort_keypoints(const KeyPoint* in, KeyPoint* out, int n, const float* K4, const float* dist5) {
  undistort_keypoints(in, out, n, K4, dist5);
}
void oo_stereo_from_rgbd(const KeyPoint* kps, const KeyPoint* kps_un, int n, const float* depth, size_t stride_floats,
                         float mbf, float* u_right, float* depth_out) {
  stereo_from_rgbd(kps, kps_un, n, depth, stride_floats, mbf, u_right, depth_out);
}
int oo_is_in_frustum(const cuorb_frustum_camera* cam, int n, const float* P, const float* N, const float* minD,
                     const float* maxD, float limit, uint8_t* in_view, float* px, float* py, float* pxr, int32_t* level,
                     float* view_cos, float* depth) {
  return is_in_frustum(*cam, n, P, N, minD, maxD, limit, in_view, px, py, pxr, level, view_cos, depth);
}
}
