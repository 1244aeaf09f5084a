// Drop-in for the frame preparation the reference does between the camera and
// ORBextractor::operator():
//   System::TrackStereo / TrackRGBD / TrackMonocular, src/System.cc:280-300, 384-386, 476:
//       needToRectify -> cv::remap(im, out, M1, M2, cv::INTER_LINEAR)
//       needToResize  -> cv::resize(im, out, settings_->newImSize())
//   Tracking::GrabImageStereo / RGBD / Monocular, src/Tracking.cc:1339-1357, 1449-1462, 1520-1533:
//       cv::cvtColor(mImGray, mImGray, COLOR_{RGB,BGR,RGBA,BGRA}2GRAY)
// One libcuorb pass does both (same integer arithmetic as OpenCV, bit-identical results); the
// fused form hands the gray plane to the extractor on the device, so the raw frame is the only
// upload.  Header-only; no CPU fallback.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/cuorb.h"
#include "ORBextractor.h"
#include "cv_compat.h"

namespace ORB_SLAM3 {

class FramePrep {
public:
  // neither rectification nor resize (System.cc:297-300 "clone" branch): gray conversion only
  FramePrep(int src_w, int src_h, int device = 0) { init(device, CUORB_PRE_NONE, src_w, src_h, src_w, src_h, nullptr, nullptr); }
  // settings_->needToResize(): newImSize
  FramePrep(int src_w, int src_h, int new_w, int new_h, int device = 0) { init(device, CUORB_PRE_RESIZE, src_w, src_h, new_w, new_h, nullptr, nullptr); }
  // settings_->needToRectify(): M1 / M2 = settings_->M1l(), M2l() (CV_32FC1, src/Settings.cc:492-510)
  FramePrep(int src_w, int src_h, const cv::Mat& M1, const cv::Mat& M2, int device = 0) {
    if (M1.type() != CV_32FC1 || M2.type() != CV_32FC1 || M1.rows != M2.rows || M1.cols != M2.cols)
      throw std::invalid_argument("FramePrep: M1/M2 must be CV_32FC1 maps of equal size");
    std::vector<float> mx((size_t)M1.rows * M1.cols), my(mx.size());
    for (int y = 0; y < M1.rows; y++) {
      std::memcpy(&mx[(size_t)y * M1.cols], M1.ptr<float>(y), sizeof(float) * M1.cols);
      std::memcpy(&my[(size_t)y * M1.cols], M2.ptr<float>(y), sizeof(float) * M1.cols);
    }
    init(device, CUORB_PRE_REMAP, src_w, src_h, M1.cols, M1.rows, mx.data(), my.data());
  }
  ~FramePrep() { cuorb_preproc_destroy(p_); }
  FramePrep(const FramePrep&) = delete;
  FramePrep& operator=(const FramePrep&) = delete;

  // imToFeed + cvtColor: raw camera frame (1, 3 or 4 channels; bRGB = Tracking::mbRGB) -> gray
  void operator()(const cv::Mat& im, bool bRGB, cv::Mat& gray) {
    gray.create(dh_, dw_, CV_8UC1);
    const int rc = cuorb_preprocess(p_, 1, im.data, im.step, 0, CUORB_MEM_HOST, format(im, bRGB), gray.data, CUORB_MEM_HOST);
    if (rc < 0) throw std::runtime_error(std::string("FramePrep (libcuorb): ") + cuorb_last_error());
  }

  // the same, fused with (*extractor)(gray, cv::Mat(), keys, descriptors, vLappingArea)
  int Extract(ORBextractor& extractor, const cv::Mat& im, bool bRGB, std::vector<cv::KeyPoint>& keys,
              cv::Mat& descriptors, std::vector<int>& vLappingArea) {
    if (im.empty()) return -1;
    const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0, lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
    int rc = cuorb_extract_batch_async_pre(extractor.handle(), p_, 1, im.data, im.step, 0, CUORB_MEM_HOST,
                                           format(im, bRGB), lap0, lap1);
    if (rc < 0) throw std::runtime_error(std::string("FramePrep::Extract (libcuorb): ") + cuorb_last_error());
    const int cap = cuorb_extractor_capacity(extractor.handle());
    std::vector<cuorb_keypoint> kp(cap);
    cv::Mat desc(cap, 32, CV_8U);
    int32_t n = 0, mono = 0;
    rc = cuorb_extract_fetch(extractor.handle(), kp.data(), desc.data, cap, &n, &mono);
    if (rc < 0) throw std::runtime_error(std::string("FramePrep::Extract (libcuorb): ") + cuorb_last_error());
    keys.resize(n);
    if (n == 0) { descriptors.release(); return mono; }
    std::memcpy((void*)keys.data(), kp.data(), sizeof(cuorb_keypoint) * (size_t)n);
    descriptors.create(n, 32, CV_8U);
    for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr<unsigned char>(i), desc.ptr<unsigned char>(i), 32);
    return mono;
  }

  int width() const { return dw_; }
  int height() const { return dh_; }
  cuorb_preproc* handle() const { return p_; }

private:
  void init(int device, cuorb_pre_mode mode, int sw, int sh, int dw, int dh, const float* mx, const float* my) {
    sw_ = sw; sh_ = sh; dw_ = dw; dh_ = dh;
    p_ = cuorb_preproc_create(device, mode, sw, sh, dw, dh, mx, my, 1);
    if (!p_) throw std::runtime_error(std::string("FramePrep (libcuorb): ") + cuorb_last_error());
  }
  cuorb_pixfmt format(const cv::Mat& im, bool bRGB) const {
    if (im.cols != sw_ || im.rows != sh_) throw std::invalid_argument("FramePrep: frame size differs from the size given at construction");
    switch (im.channels()) {                       // src/Tracking.cc:1339-1357
      case 1: return CUORB_PIX_GRAY8;
      case 3: return bRGB ? CUORB_PIX_RGB8 : CUORB_PIX_BGR8;
      case 4: return bRGB ? CUORB_PIX_RGBA8 : CUORB_PIX_BGRA8;
      default: throw std::invalid_argument("FramePrep: frame must have 1, 3 or 4 channels");
    }
  }
  cuorb_preproc* p_ = nullptr;
  int sw_ = 0, sh_ = 0, dw_ = 0, dh_ = 0;
};

}  // namespace ORB_SLAM3
