// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stand-in for Sophus SE3 / Sim3 (the vendored
// Thirdparty/Sophus needs Eigen, absent here).  Rotation is held as a matrix (the real SO3 holds a
// unit quaternion and rotates through it): results are deterministic, not Sophus-bit-identical —
// parity of the pose algebra is NOT claimed, only of everything the reference does with the
// projected points (see ref_capi_match.cpp).
#pragma once
#include <Eigen/Core>

namespace Sophus {

template <typename T> class SO3 {
public:
  // the cross-product matrix [w]x
  static Eigen::Matrix<T, 3, 3> hat(const Eigen::Matrix<T, 3, 1>& w) {
    Eigen::Matrix<T, 3, 3> m;
    m(0, 1) = -w(2); m(0, 2) = w(1); m(1, 0) = w(2); m(1, 2) = -w(0); m(2, 0) = -w(1); m(2, 1) = w(0);
    return m;
  }
};
typedef SO3<float> SO3f;

template <typename T> class SE3 {
public:
  typedef Eigen::Matrix<T, 3, 3> Mat3;
  typedef Eigen::Matrix<T, 3, 1> Vec3;
  SE3() : R_(Mat3::Identity()) {}
  SE3(const Mat3& R, const Vec3& t) : R_(R), t_(t) {}
  const Mat3& rotationMatrix() const { return R_; }
  const Vec3& translation() const { return t_; }
  SE3  inverse() const { Mat3 Rt = R_.transpose(); return SE3(Rt, -(Rt * t_)); }
  Vec3 operator*(const Vec3& p) const { return R_ * p + t_; }
  SE3  operator*(const SE3& o) const { return SE3(R_ * o.R_, R_ * o.t_ + t_); }
private:
  Mat3 R_;
  Vec3 t_;
};
typedef SE3<float>  SE3f;
typedef SE3<double> SE3d;

template <typename T> class Sim3 {
public:
  typedef Eigen::Matrix<T, 3, 3> Mat3;
  typedef Eigen::Matrix<T, 3, 1> Vec3;
  Sim3() : s_(1), R_(Mat3::Identity()) {}
  Sim3(T s, const Mat3& R, const Vec3& t) : s_(s), R_(R), t_(t) {}
  T           scale() const { return s_; }
  const Mat3& rotationMatrix() const { return R_; }
  const Vec3& translation() const { return t_; }
  Sim3 inverse() const { Mat3 Rt = R_.transpose(); T is = T(1) / s_; return Sim3(is, Rt, -((Rt * t_) * is)); }
  Vec3 operator*(const Vec3& p) const { return (R_ * p) * s_ + t_; }
private:
  T    s_;
  Mat3 R_;
  Vec3 t_;
};
typedef Sim3<float>  Sim3f;
typedef Sim3<double> Sim3d;

}  // namespace Sophus
