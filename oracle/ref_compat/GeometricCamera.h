// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stub of include/CameraModels/GeometricCamera.h and
// Pinhole.h: the virtual interface src/ORBmatcher.cc and the lifted Frame functions call.
// Pinhole::project is the reference's own body (lifted from src/CameraModels/Pinhole.cpp:55-76
// at build time); epipolarConstrain is delegated to the harness.
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
  // harness hook: the epipolar test needs K^-T [t]x R K^-1 (matrix inverse, SO3::hat) which the
  // Eigen stand-in does not carry; a harness that exercises SearchForTriangulation installs one
  bool (*epipolar_hook)(const cv::KeyPoint&, const cv::KeyPoint&, const Eigen::Matrix3f&,
                        const Eigen::Vector3f&, float, float, void*) = nullptr;
  void* epipolar_ctx = nullptr;
  bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2,
                         const Eigen::Matrix3f& R12, const Eigen::Vector3f& t12,
                         const float sigmaLevel, const float unc) {
    if (!epipolar_hook) REF_COMPAT_UNSUPPORTED("Pinhole::epipolarConstrain without a harness hook");
    return epipolar_hook(kp1, kp2, R12, t12, sigmaLevel, unc, epipolar_ctx);
  }
};

}  // namespace ORB_SLAM3
