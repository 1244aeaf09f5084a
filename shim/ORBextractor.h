// Drop-in replacement for the reference's include/ORBextractor.h (class ORB_SLAM3::ORBextractor,
// include/ORBextractor.h:41-119): same constructor, same operator() signature / return value /
// output conventions, same getters, same public mvImagePyramid member.  All compute is done by
// libcuorb (include/cuorb.h); there is no CPU fallback — construction throws when no CUDA device
// is usable.  Header-only on purpose: the maintainer replaces one header and links -lcuorb.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/cuorb.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

class ORBextractor {
public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  // include/ORBextractor.h:48.  max_width/max_height size the device arena; images larger than
  // that make operator() fail (they are the camera resolution in the reference's use).
  ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_,
               int max_width = 1920, int max_height = 1200, int device = 0)
      : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
        minThFAST(minThFAST_) {
    cuorb_config cfg;
    cfg.nfeatures = nfeatures_; cfg.scale_factor = scaleFactor_; cfg.nlevels = nlevels_;
    cfg.ini_th_fast = iniThFAST_; cfg.min_th_fast = minThFAST_;
    cfg.max_width = max_width; cfg.max_height = max_height; cfg.max_batch = 1; cfg.device = device;
    h_ = cuorb_extractor_create(&cfg);
    if (!h_) throw std::runtime_error(std::string("ORBextractor (libcuorb): ") + cuorb_last_error());
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    cuorb_extractor_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data(), mnFeaturesPerLevel.data());
    mvImagePyramid.resize(nlevels);
    capacity_ = cuorb_extractor_capacity(h_);
    kp_buf_.resize(capacity_);
  }
  ~ORBextractor() { cuorb_extractor_destroy(h_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // include/ORBextractor.h:55-61, src/ORBextractor.cc:1052-1138.  Returns monoIndex; -1 when the
  // image is empty (outputs untouched, as in the reference); mask ignored; _keypoints replaced;
  // _descriptors (re)created n x 32 CV_8U, or released when n == 0.
  int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                 cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
    if (_image.empty()) return -1;
    cv::Mat image = _image.getMat();
    if (image.type() != CV_8UC1) CUORB_SHIM_FAIL(std::invalid_argument, "ORBextractor: image must be CV_8UC1", -1);
    const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0;
    const int lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
    desc_buf_.create(capacity_, 32, CV_8U);
    int n = 0;
    const int mono = cuorb_extract(h_, image.data, image.cols, image.rows, image.step, lap0, lap1,
                                   kp_buf_.data(), desc_buf_.data, capacity_, &n);
    if (mono < 0) CUORB_SHIM_FAIL(std::runtime_error, std::string("ORBextractor (libcuorb): ") + cuorb_last_error(), -1);
    pyramid_stale_ = true;
    static_assert(sizeof(cv::KeyPoint) == sizeof(cuorb_keypoint), "cv::KeyPoint must be 7 x 4 bytes");
    _keypoints.resize(n);
    if (n == 0) { _descriptors.release(); return mono; }
    std::memcpy((void*)_keypoints.data(), kp_buf_.data(), sizeof(cuorb_keypoint) * (size_t)n);
    _descriptors.create(n, 32, CV_8U);
    cv::Mat out = _descriptors.getMat();   // cv::OutputArray has no ptr<>(): go through the Mat header it wraps
    for (int i = 0; i < n; i++) std::memcpy(out.ptr<unsigned char>(i), desc_buf_.ptr<unsigned char>(i), 32);
    return mono;
  }

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return (float)scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  // include/ORBextractor.h:87.  The pyramid lives on the device (stereo matching reads it there);
  // host copies are made on demand by SyncImagePyramid() for code that still reads this member.
  std::vector<cv::Mat> mvImagePyramid;
  void SyncImagePyramid() {
    if (!pyramid_stale_) return;
    for (int l = 0; l < nlevels; l++) {
      const uint8_t* dev; int w, h; size_t st;
      if (cuorb_extractor_pyramid(h_, 0, l, &dev, &w, &h, &st) != CUORB_OK) return;
      mvImagePyramid[l].create(h, w, CV_8UC1);
      cuorb_extractor_download_plane(h_, 0, l, 1, mvImagePyramid[l].data, (size_t)w * h, nullptr, nullptr);
    }
    pyramid_stale_ = false;
  }

  cuorb_extractor* handle() const { return h_; }  // for ORBmatcher / Frame::ComputeStereoMatches
  void InvalidateImagePyramid() { pyramid_stale_ = true; }   // after an extraction that bypassed operator()

protected:
  int    nfeatures;
  double scaleFactor;
  int    nlevels;
  int    iniThFAST;
  int    minThFAST;
  std::vector<int>   mnFeaturesPerLevel;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

private:
  cuorb_extractor*            h_ = nullptr;
  int                         capacity_ = 0;
  std::vector<cuorb_keypoint> kp_buf_;
  cv::Mat                     desc_buf_;
  bool                        pyramid_stale_ = true;
};

}  // namespace ORB_SLAM3
