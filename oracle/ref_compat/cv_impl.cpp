// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// The OpenCV functions the reference's path calls, forwarded from the ref_compat cv:: layer to
// the restated primitives of oracle/orb_oracle.cpp (pinned bit-for-bit against the real cv2
// 4.13.0 by tests/test_oracle_golden.py).  Argument combinations the primitives do not cover
// abort instead of guessing.
#include <opencv2/core.hpp>
#include <opencv2/features2d.hpp>
#include <opencv2/imgproc.hpp>

#include "../orb_oracle.hpp"

namespace {
orb_oracle::Img view(const cv::Mat& m) {
  orb_oracle::Img im;
  im.data = const_cast<uint8_t*>(m.data); im.cols = m.cols; im.rows = m.rows; im.step = m.step;
  return im;
}
void need_8uc1(const cv::Mat& m, const char* who) {
  if (m.type() != CV_8UC1) REF_COMPAT_UNSUPPORTED(who);
}
}  // namespace

namespace cv {

// src/ORBextractor.cc:1150 — cv::resize(level l-1 ROI, level l ROI, sz, 0, 0, INTER_LINEAR)
void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation) {
  Mat src = src_.getMat();
  need_8uc1(src, "cv::resize on a type other than CV_8UC1");
  if (interpolation != INTER_LINEAR || fx != 0 || fy != 0 || dsize.width <= 0 || dsize.height <= 0)
    REF_COMPAT_UNSUPPORTED("cv::resize with anything but (dsize, 0, 0, INTER_LINEAR)");
  // exact integer decimation by 2 takes OpenCV's INTER_AREA fast path: not restated
  if (src.cols == 2 * dsize.width && src.rows == 2 * dsize.height)
    REF_COMPAT_UNSUPPORTED("cv::resize exact 2x decimation (INTER_AREA path)");
  dst_.create(dsize, src.type());  // no-op for the pyramid ROI: writes go through to `temp`
  Mat dst = dst_.getMat();
  orb_oracle::Img s = view(src), d = view(dst);
  orb_oracle::resize_linear_8u(s, d);
}

// src/ORBextractor.cc:1152-1170 — equal borders, BORDER_REFLECT_101 (+ BORDER_ISOLATED when the
// source is the pyramid ROI, i.e. pixels outside the ROI are NOT used)
void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right,
                    int borderType) {
  Mat src = src_.getMat();
  need_8uc1(src, "cv::copyMakeBorder on a type other than CV_8UC1");
  const bool isolated = (borderType & BORDER_ISOLATED) != 0;
  if ((borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101 || top != bottom || top != left ||
      top != right)
    REF_COMPAT_UNSUPPORTED("cv::copyMakeBorder other than equal REFLECT_101 borders");
  // without BORDER_ISOLATED the real function extends a sub-matrix with its parent's pixels
  if (!isolated && src.isSubmatrix())
    REF_COMPAT_UNSUPPORTED("cv::copyMakeBorder of a sub-matrix without BORDER_ISOLATED");
  dst_.create(src.rows + top + bottom, src.cols + left + right, src.type());
  Mat dst = dst_.getMat();
  orb_oracle::Img s = view(src), d = view(dst);
  orb_oracle::copy_make_border_reflect101(s, d, top);
}

// src/ORBextractor.cc:1104 — GaussianBlur(m, m, Size(7,7), 2, 2, BORDER_REFLECT_101), in place
void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY,
                  int borderType) {
  Mat src = src_.getMat();
  need_8uc1(src, "cv::GaussianBlur on a type other than CV_8UC1");
  if (ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 ||
      (borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101)
    REF_COMPAT_UNSUPPORTED("cv::GaussianBlur other than 7x7, sigma 2, REFLECT_101");
  if (src.isSubmatrix() && !(borderType & BORDER_ISOLATED))
    REF_COMPAT_UNSUPPORTED("cv::GaussianBlur of a sub-matrix without BORDER_ISOLATED");
  dst_.create(src.rows, src.cols, src.type());
  Mat dst = dst_.getMat();
  orb_oracle::Img s = view(src), d = view(dst);
  orb_oracle::gaussian_blur_7x7_s2(s, d);  // row pass completes into scratch first: in-place safe
}

// src/ORBextractor.cc:800,821 — cv::FAST(cell ROI, keys, threshold, true), TYPE_9_16
void FAST(InputArray image_, std::vector<KeyPoint>& keypoints, int threshold, bool nonmax) {
  Mat image = image_.getMat();
  need_8uc1(image, "cv::FAST on a type other than CV_8UC1");
  std::vector<orb_oracle::KeyPoint> k;
  orb_oracle::fast9_16(view(image), k, threshold, nonmax);
  keypoints.resize(k.size());
  static_assert(sizeof(orb_oracle::KeyPoint) == sizeof(KeyPoint), "KeyPoint layout");
  if (!k.empty()) std::memcpy((void*)keypoints.data(), k.data(), k.size() * sizeof(KeyPoint));
}

float fastAtan2(float y, float x) { return orb_oracle::fast_atan2(y, x); }

// src/Frame.cc:1060 — cv::norm(IL, IR, NORM_L1) on two 11x11 CV_8U windows: sum |a-b|, as double
double norm(InputArray a_, InputArray b_, int normType) {
  Mat a = a_.getMat(), b = b_.getMat();
  need_8uc1(a, "cv::norm on a type other than CV_8UC1");
  need_8uc1(b, "cv::norm on a type other than CV_8UC1");
  if (normType != NORM_L1 || a.rows != b.rows || a.cols != b.cols)
    REF_COMPAT_UNSUPPORTED("cv::norm other than NORM_L1 of equal-size 8-bit windows");
  long s = 0;
  for (int y = 0; y < a.rows; y++) {
    const uchar *p = a.ptr(y), *q = b.ptr(y);
    for (int x = 0; x < a.cols; x++) s += std::abs((int)p[x] - (int)q[x]);
  }
  return (double)s;
}

// only reachable from the dead ComputeKeyPointsOld; keeps the strongest npoints (ties kept)
void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int npoints) {
  if (npoints >= 0 && keypoints.size() > (size_t)npoints) {
    if (npoints == 0) { keypoints.clear(); return; }
    std::nth_element(keypoints.begin(), keypoints.begin() + npoints - 1, keypoints.end(),
                     [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
    float ambiguous = keypoints[npoints - 1].response;
    auto  e = std::partition(keypoints.begin() + npoints, keypoints.end(),
                             [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
    keypoints.resize(e - keypoints.begin());
  }
}

}  // namespace cv

// cv::BFMatcher::knnMatch(query, train, matches, 2): rows in (distance, trainIdx) order, fewer when the train set is small
namespace orb_oracle { void knn_match2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist); }
void cv::BFMatcher::knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch>>& matches, int k) const {
  if (k != 2) REF_COMPAT_UNSUPPORTED("BFMatcher::knnMatch with k != 2");
  if (q.type() != CV_8UC1 || (q.rows && q.cols != 32) || (t.rows && t.cols != 32)) REF_COMPAT_UNSUPPORTED("knnMatch: not n x 32 CV_8U");
  const int nq = q.rows, nt = t.rows;
  std::vector<uint8_t> qb((size_t)std::max(nq, 1) * 32), tb((size_t)std::max(nt, 1) * 32);
  for (int i = 0; i < nq; i++) memcpy(&qb[(size_t)i * 32], q.ptr<uint8_t>(i), 32);
  for (int i = 0; i < nt; i++) memcpy(&tb[(size_t)i * 32], t.ptr<uint8_t>(i), 32);
  std::vector<int32_t> idx((size_t)std::max(nq, 1) * 2, -1), dist((size_t)std::max(nq, 1) * 2, -1);
  if (nq && nt) orb_oracle::knn_match2(qb.data(), nq, tb.data(), nt, idx.data(), dist.data());
  matches.assign(nq, {});
  for (int i = 0; i < nq; i++)
    for (int j = 0; j < 2; j++)
      if (idx[2 * i + j] >= 0) {
        DMatch m;
        m.queryIdx = i; m.trainIdx = idx[2 * i + j]; m.imgIdx = 0; m.distance = (float)dist[2 * i + j];
        matches[i].push_back(m);
      }
}
