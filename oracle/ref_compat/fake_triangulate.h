// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  A deterministic stand-in for KannalaBrandt8::TriangulateMatches (the camera
// model is out of scope): the same function is installed on the reference side (oracle/ref_compat/ref_capi_rig.cpp)
// and on the shim side (tests/native/fisheye_check.cpp), so everything AROUND the triangulation can be compared.
#pragma once
#include <Eigen/Core>
#include <opencv2/core.hpp>

inline float fake_triangulate(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, float sigma1, float sigma2, Eigen::Vector3f& p3D) {
  float depth = (kp1.pt.x - kp2.pt.x) * 0.01f + sigma1 * 0.001f - sigma2 * 0.002f;
  if (((int)(kp1.pt.y + kp2.pt.y)) % 5 == 0) depth = -1.f;     // some matches are rejected by the depth test
  p3D = Eigen::Vector3f(kp1.pt.x, kp2.pt.y, depth);
  return depth;
}
