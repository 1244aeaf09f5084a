// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stub of include/CameraModels/GeometricCamera.h and
// Pinhole.h: the virtual interface src/ORBmatcher.cc and the lifted Frame functions call.
// Pinhole::project and Pinhole::epipolarConstrain are the reference's own bodies (lifted from
// src/CameraModels/Pinhole.cpp:55-76 and :143-172 at build time).
#pragma once
#include <vector>
#include <Eigen/Core>
#include <opencv2/core.hpp>
#include <sophus/se3.hpp>

namespace ORB_SLAM3 {

class GeometricCamera {
public:
  GeometricCamera() {}
  GeometricCamera(const std::vector<float>& p) : mvParameters(p) {}
  virtual ~GeometricCamera() {}
  virtual cv::Point2f     project(const cv::Point3f& p3D)     = 0;
  virtual Eigen::Vector2d project(const Eigen::Vector3d& v3D) = 0;
  virtual Eigen::Vector2f project(const Eigen::Vector3f& v3D) = 0;
  virtual bool epipolarConstrain(GeometricCamera* otherCamera, const cv::KeyPoint& kp1,
                                 const cv::KeyPoint& kp2, const Eigen::Matrix3f& R12,
                                 const Eigen::Vector3f& t12, const float sigmaLevel,
                                 const float unc) = 0;
  virtual Eigen::Matrix3f toK_() { return Eigen::Matrix3f::Identity(); }
  float getParameter(const int i) { return mvParameters[i]; }
protected:
  std::vector<float> mvParameters;  // [fx, fy, cx, cy]
};

class Pinhole : public GeometricCamera {
public:
  Pinhole(const std::vector<float>& p) : GeometricCamera(p) {}
  cv::Point2f     project(const cv::Point3f& p3D);
  Eigen::Vector2d project(const Eigen::Vector3d& v3D);
  Eigen::Vector2f project(const Eigen::Vector3f& v3D);
  // K as an Eigen matrix (Pinhole.cpp:137-141 fills it with the comma initialiser, absent from the stand-in)
  Eigen::Matrix3f toK_() {
    Eigen::Matrix3f K;
    K(0, 0) = mvParameters[0]; K(0, 2) = mvParameters[2]; K(1, 1) = mvParameters[1]; K(1, 2) = mvParameters[3];
    K(2, 2) = 1.f;
    return K;
  }
  // the reference's own body (Pinhole.cpp:143-172, lifted); its F12 = K1^-T [t12]x R12 K2^-1 runs on the
  // Eigen / Sophus stand-ins (deterministic, not Eigen-bit-identical), the point-to-line test after it is
  // the reference's arithmetic
  bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                         const Eigen::Matrix3f& R12, const Eigen::Vector3f& t12, const float sigmaLevel,
                         const float unc);
};

// Stub of include/CameraModels/KannalaBrandt8.h: Frame::ComputeStereoFishEyeMatches static_casts mpCamera to this
// class and calls TriangulateMatches (src/Frame.cc:1335-1344).  The camera model itself is out of scope (SURVEY.md
// §2); a harness installs a deterministic stand-in for the triangulation and checks everything around it.
class KannalaBrandt8 : public GeometricCamera {
public:
  KannalaBrandt8() {}
  cv::Point2f     project(const cv::Point3f& p) override { return cv::Point2f(p.x, p.y); }
  Eigen::Vector2d project(const Eigen::Vector3d& v) override { return Eigen::Vector2d(v[0], v[1]); }
  Eigen::Vector2f project(const Eigen::Vector3f& v) override { return Eigen::Vector2f(v[0], v[1]); }
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&, const Eigen::Matrix3f&,
                         const Eigen::Vector3f&, const float, const float) override { return false; }
  float (*triangulate_hook)(const cv::KeyPoint&, const cv::KeyPoint&, float, float, Eigen::Vector3f&) = nullptr;
  float TriangulateMatches(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const Eigen::Matrix3f&,
                           const Eigen::Vector3f&, const float sigmaLevel, const float unc, Eigen::Vector3f& p3D) {
    return triangulate_hook(kp1, kp2, sigmaLevel, unc, p3D);
  }
};

}  // namespace ORB_SLAM3
