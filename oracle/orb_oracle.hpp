// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// CPU oracle: a plain C++ restatement of the reference's ORB front-end path
// (elkuno213/orbslam3_cpp_fork: src/ORBextractor.cc, src/Frame.cc, src/ORBmatcher.cc) with the
// OpenCV primitives it calls (cv::resize INTER_LINEAR, copyMakeBorder REFLECT_101, cv::FAST,
// cv::GaussianBlur 7x7 sigma 2, cv::fastAtan2, cvRound, cv::norm L1) restated from OpenCV's
// published algorithms and pinned against the real library (python cv2 4.13.0) by
// tests/test_oracle_cv2.py and the committed golden vectors in tests/golden/.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this.  libcuorb never links or calls it.
//
// Parity status: pinned.  The reference's own tests hold no golden vectors for this path, so
// (1) the OpenCV primitives are pinned to outputs of the real OpenCV 4.13.0 through cv2 (tests/golden/), and
// (2) everything above them is pinned to the reference's OWN code: src/ORBextractor.cc and src/ORBmatcher.cc
// compiled unmodified, plus bodies lifted verbatim from src/Frame.cc, src/KeyFrame.cc, src/MapPoint.cc and
// Pinhole.cpp, over the stand-in headers of oracle/ref_compat (oracle/_ref; tests/test_oracle_vs_ref*.py).
// Behaviour is defined as: OpenCV 4.13.0, GCC 13 libstdc++ (std::list / std::sort tie order in
// DistributeOctTree), glibc cosf/sinf, compiled -O2 -ffp-contract=off (no FMA contraction).
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>

namespace orb_oracle {

struct KeyPoint {  // cv::KeyPoint layout
  float   x = 0, y = 0, size = 0, angle = -1, response = 0;
  int32_t octave = 0, class_id = -1;
};

// 8-bit single-channel image view (ROI of a larger buffer allowed, like cv::Mat)
struct Img {
  uint8_t* data = nullptr;
  int      cols = 0, rows = 0;
  size_t   step = 0;
  uint8_t*       ptr(int y) { return data + (size_t)y * step; }
  const uint8_t* ptr(int y) const { return data + (size_t)y * step; }
};

// ---- restated OpenCV primitives ----
int   cv_round(float v);
int   cv_round(double v);
void  resize_linear_8u(const Img& src, Img& dst);
void  copy_make_border_reflect101(const Img& src, Img& dst_whole, int border);
void  fast9_16(const Img& roi, std::vector<KeyPoint>& out, int threshold, bool nonmax);
void  gaussian_blur_7x7_s2(const Img& src, Img& dst);
float fast_atan2(float y, float x);
int   fast_corner_score(const uint8_t* p, size_t step);  // max(A,B)-1, threshold independent

struct Pyramid {
  std::vector<std::vector<uint8_t>> storage;  // bordered planes ("temp")
  std::vector<Img>                  level;    // ROI views (mvImagePyramid)
  std::vector<Img>                  whole;
};

class ORBextractor {
public:
  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
  int operator()(const Img& image, std::vector<KeyPoint>& keypoints,
                 std::vector<uint8_t>& descriptors, const int lapping[2]);

  int    nfeatures;
  double scaleFactor;
  int    nlevels, iniThFAST, minThFAST;
  std::vector<int>   mnFeaturesPerLevel, umax;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  Pyramid            pyr;  // mvImagePyramid

  // stage taps for parity tests
  std::vector<std::vector<KeyPoint>>  dbg_candidates;  // vToDistributeKeys per level
  std::vector<std::vector<KeyPoint>>  dbg_level_keys;  // allKeypoints per level (with angle)
  std::vector<std::vector<uint8_t>>   dbg_blurred;     // blurred level (tight w*h), empty if skipped

  void ComputePyramid(const Img& image);
  void ComputeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& allKeypoints);
  std::vector<KeyPoint> DistributeOctTree(const std::vector<KeyPoint>& vToDistributeKeys, int minX,
                                          int maxX, int minY, int maxY, int N, int level);
};

int descriptor_distance(const uint8_t* a, const uint8_t* b);

// Frame::ComputeStereoMatches (src/Frame.cc:944-1113)
struct StereoDebug {
  std::vector<int32_t> best_idx_r, best_dist, sad;
};
void compute_stereo_matches(const ORBextractor& L, const ORBextractor& R,
                            const std::vector<KeyPoint>& keysL, const uint8_t* descL,
                            const std::vector<KeyPoint>& keysR, const uint8_t* descR, float mb,
                            float mbf, std::vector<float>& uRight, std::vector<float>& depth,
                            StereoDebug* dbg);

// The Frame members the projection searches read
struct FrameView {
  int                  N = 0;
  const KeyPoint*      keysUn = nullptr;
  const uint8_t*       desc = nullptr;
  const float*         uRight = nullptr;
  std::vector<uint8_t> bound_with_obs;  // mutable copy: mvpMapPoints[i] && Observations()>0
  float minX = 0, minY = 0, maxX = 0, maxY = 0, gridWInv = 0, gridHInv = 0, mb = 0, mbf = 0;
  std::vector<float>   scaleFactors;
  std::vector<std::vector<int>> grid;   // FRAME_GRID_COLS*FRAME_GRID_ROWS cells, [ix*48+iy]
  void assign_features_to_grid();       // src/Frame.cc:524-553, 867-877
  std::vector<int> features_in_area(float x, float y, float r, int minLevel, int maxLevel) const;
};

int search_by_projection_mappoints(FrameView& F, int M, const uint8_t* mpDesc, const float* projX,
                                   const float* projY, const float* projXR, const int32_t* level,
                                   const float* viewCos, const float* depth, const uint8_t* flags,
                                   float th, bool bFarPoints, float thFarPoints, float nnratio,
                                   int32_t* bestIdxOut);
int search_by_projection_frame(FrameView& Cur, int nLast, const uint8_t* lastDesc, const float* u,
                               const float* v, const float* invz, const int32_t* lastOctave,
                               const float* lastAngle, const uint8_t* flags, float th,
                               bool bForward, bool bBackward, bool checkOri, int32_t* bestIdxOut,
                               int32_t* rotBinOut, int32_t* keepBinsOut);

}  // namespace orb_oracle
