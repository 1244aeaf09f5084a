// Frame::isInFrustum for one MapPoint, shared by k_is_in_frustum (frameglue.cu) and the resident local map
// (k_localmap_frustum, project.cu).
#pragma once
#include "common.cuh"

namespace cuorb {

// Frame::isInFrustum (src/Frame.cc:644-707, Nleft == -1) + MapPoint::PredictScale (src/MapPoint.cc:574-589) +
// Pinhole::project (Pinhole.cpp:70-76).  Every float operation is an explicit round-to-nearest intrinsic in the
// reference's expression order (size-3 sums associate as Eigen's unrolled reduction does: x0 + (x1 + x2)), so nvcc
// cannot contract anything into an FMA.
__device__ __forceinline__ float dot3(float a0, float a1, float a2, float b0, float b1, float b2) {
  return __fadd_rn(__fmul_rn(a0, b0), __fadd_rn(__fmul_rn(a1, b1), __fmul_rn(a2, b2)));
}

struct FrustumOut { float px, py, pxr, vc, dz; int lv; };

// returns mbTrackInView; o holds mTrackProjX / Y / XR, mnTrackScaleLevel, mTrackViewCos, mTrackDepth as the
// reference leaves them (unset fields keep the caller's initial values: -1 / -1 / -1 / -1 / 0 / 0)
__device__ __forceinline__ bool frustum_point(const cuorb_frustum_camera& c, float P0, float P1, float P2, float N0, float N1,
                                              float N2, float minD, float mx, float limit, FrustumOut& o) {
  o.px = -1.f; o.py = -1.f; o.pxr = -1.f; o.vc = 0.f; o.dz = 0.f; o.lv = -1;
  const float Pc0 = __fadd_rn(dot3(c.rcw[0], c.rcw[1], c.rcw[2], P0, P1, P2), c.tcw[0]);
  const float Pc1 = __fadd_rn(dot3(c.rcw[3], c.rcw[4], c.rcw[5], P0, P1, P2), c.tcw[1]);
  const float Pc2 = __fadd_rn(dot3(c.rcw[6], c.rcw[7], c.rcw[8], P0, P1, P2), c.tcw[2]);
  const float Pc_dist = __fsqrt_rn(dot3(Pc0, Pc1, Pc2, Pc0, Pc1, Pc2));
  const float invz = __fdiv_rn(1.0f, Pc2);
  if (Pc2 < 0.0f) return false;
  const float u = __fadd_rn(__fdiv_rn(__fmul_rn(c.fx, Pc0), Pc2), c.cx);
  const float v = __fadd_rn(__fdiv_rn(__fmul_rn(c.fy, Pc1), Pc2), c.cy);
  if ((u < c.min_x || u > c.max_x) || (v < c.min_y || v > c.max_y)) return false;
  o.px = u; o.py = v;
  const float maxDistance = __fmul_rn(1.2f, mx), minDistance = __fmul_rn(0.8f, minD);
  const float O0 = __fsub_rn(P0, c.ow[0]), O1 = __fsub_rn(P1, c.ow[1]), O2 = __fsub_rn(P2, c.ow[2]);
  const float dist = __fsqrt_rn(dot3(O0, O1, O2, O0, O1, O2));
  if (dist < minDistance || dist > maxDistance) return false;
  const float viewCos = __fdiv_rn(dot3(O0, O1, O2, N0, N1, N2), dist);
  if (viewCos < limit) return false;
  const float ratio = __fdiv_rn(mx, dist);
  // std::log(float) = glibc logf (correctly rounded in practice): double log, rounded once
  const float lg = (float)log((double)ratio);
  int nScale = (int)ceilf(__fdiv_rn(lg, c.log_scale_factor));
  nScale = nScale < 0 ? 0 : (nScale >= c.n_levels ? c.n_levels - 1 : nScale);
  o.pxr = __fsub_rn(u, __fmul_rn(c.mbf, invz));
  o.dz = Pc_dist; o.lv = nScale; o.vc = viewCos;
  return true;
}

}  // namespace cuorb
