// cv::undistortPoints for one point, shared by k_undistort (frameglue.cu) and the device-side Frame build of the
// resident local-map search (k_frame_build, project.cu).
#pragma once
#include "common.cuh"

namespace cuorb {

struct UndistortArgs {
  double fx, fy, cx, cy, ifx, ify;
  double k[12];
};

// K4 = fx, fy, cx, cy; dist5 = k1, k2, p1, p2, k3 (the float entries of mK / mDistCoef widened to double, as
// cv::undistortPoints does on entry)
inline UndistortArgs undistort_args(const float* K4, const float* dist5) {
  UndistortArgs a;
  a.fx = K4[0]; a.fy = K4[1]; a.cx = K4[2]; a.cy = K4[3]; a.ifx = 1. / a.fx; a.ify = 1. / a.fy;
  for (int i = 0; i < 12; i++) a.k[i] = i < 5 ? (double)dist5[i] : 0.0;
  return a;
}

// Frame::UndistortKeyPoints (src/Frame.cc:886-913) = cv::undistortPoints(mat, mat, K, D, cv::Mat(), K): calib3d's
// cvUndistortPointsInternal in doubles, 5 fixed iterations, projection by P = K, result rounded to float.  Explicit
// round-to-nearest double intrinsics in OpenCV's operation order, so no FMA contraction can change a bit.
__device__ __forceinline__ void undistort_point(const UndistortArgs& a, float& px, float& py) {
  double x = (double)px, y = (double)py;
  const double u = x, v = y;
  x = __dmul_rn(__dsub_rn(x, a.cx), a.ifx); y = __dmul_rn(__dsub_rn(y, a.cy), a.ify);
  const double x0 = x, y0 = y;
  const double* k = a.k;
  for (int j = 0; j < 5; j++) {
    const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
    const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[7], r2), k[6]), r2), k[5]), r2));
    const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[4], r2), k[1]), r2), k[0]), r2));
    const double icdist = __ddiv_rn(num, den);
    if (icdist < 0) { x = __dmul_rn(__dsub_rn(u, a.cx), a.ifx); y = __dmul_rn(__dsub_rn(v, a.cy), a.ify); break; }
    // deltaX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2   (left to right)
    double dX = __dmul_rn(__dmul_rn(__dmul_rn(2.0, k[2]), x), y);
    dX = __dadd_rn(dX, __dmul_rn(k[3], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
    dX = __dadd_rn(dX, __dmul_rn(k[8], r2));
    dX = __dadd_rn(dX, __dmul_rn(__dmul_rn(k[9], r2), r2));
    // deltaY = k2*(r2 + 2*y*y) + 2*k3*x*y + k10*r2 + k11*r2*r2
    double dY = __dmul_rn(k[2], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y)));
    dY = __dadd_rn(dY, __dmul_rn(__dmul_rn(__dmul_rn(2.0, k[3]), x), y));
    dY = __dadd_rn(dY, __dmul_rn(k[10], r2));
    dY = __dadd_rn(dY, __dmul_rn(__dmul_rn(k[11], r2), r2));
    x = __dmul_rn(__dsub_rn(x0, dX), icdist);
    y = __dmul_rn(__dsub_rn(y0, dY), icdist);
  }
  const double xx = __dadd_rn(__dadd_rn(__dmul_rn(a.fx, x), __dmul_rn(0.0, y)), a.cx);
  const double yy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(a.fy, y)), a.cy);
  px = (float)xx;   // ww = 1 / (0*x + 0*y + 1) = 1
  py = (float)yy;
}

}  // namespace cuorb
