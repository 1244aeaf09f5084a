// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See orb_oracle.hpp for scope and parity status.
// Extractor half of the oracle: restated OpenCV primitives + ORBextractor.
// Reference citations are relative to the reference checkout (elkuno213/orbslam3_cpp_fork).
#include "orb_oracle.hpp"

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>

#include "../include/cuorb_pattern.h"

namespace orb_oracle {

// ------------------------------------------------------------------------------------------
// OpenCV primitives, restated
// ------------------------------------------------------------------------------------------

// cvRound: SSE cvtss2si / cvtsd2si under the default MXCSR mode = round half to even.
int cv_round(float v) { return (int)lrintf(v); }
int cv_round(double v) { return (int)lrint(v); }

static inline short sat_short(float v) {
  int i = cv_round(v);
  return (short)std::min(std::max(i, -32768), 32767);
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1 (called at src/ORBextractor.cc:1150).
// OpenCV imgproc/resize.cpp generic path: 11-bit fixed-point coefficients
// (INTER_RESIZE_COEF_BITS = 11), HResizeLinear<uchar,int,short>, then the 8u vertical pass
//   dst = ( ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2 ) >> 2 .
void resize_linear_8u(const Img& src, Img& dst) {
  const int sw = src.cols, sh = src.rows, dw = dst.cols, dh = dst.rows;
  const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
  const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;

  std::vector<int>   xofs(dw), yofs(dh);
  std::vector<short> ialpha(2 * (size_t)dw), ibeta(2 * (size_t)dh);
  for (int dx = 0; dx < dw; dx++) {
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int   sx = (int)std::floor(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    xofs[dx]           = sx;
    ialpha[2 * dx]     = sat_short((1.f - fx) * 2048.f);
    ialpha[2 * dx + 1] = sat_short(fx * 2048.f);
  }
  for (int dy = 0; dy < dh; dy++) {
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int   sy = (int)std::floor(fy);
    fy -= sy;
    yofs[dy]          = sy;
    ibeta[2 * dy]     = sat_short((1.f - fy) * 2048.f);
    ibeta[2 * dy + 1] = sat_short(fy * 2048.f);
  }
  auto clip = [](int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; };
  std::vector<int> row0(dw), row1(dw);
  auto hresize = [&](int sy, std::vector<int>& D) {
    const uint8_t* S = src.ptr(sy);
    for (int dx = 0; dx < dw; dx++) {
      int sx = xofs[dx];
      D[dx]  = (sx + 1 < sw) ? S[sx] * ialpha[2 * dx] + S[sx + 1] * ialpha[2 * dx + 1]
                             : S[sx] * 2048;
    }
  };
  for (int dy = 0; dy < dh; dy++) {
    hresize(clip(yofs[dy], 0, sh), row0);
    hresize(clip(yofs[dy] + 1, 0, sh), row1);
    const int b0 = ibeta[2 * dy], b1 = ibeta[2 * dy + 1];
    uint8_t*  d  = dst.ptr(dy);
    for (int x = 0; x < dw; x++)
      d[x] = (uint8_t)((((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2);
  }
}

static inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
  return i;
}

// cv::copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101) (src/ORBextractor.cc:1152-1170).
// dst_whole is (cols+2b) x (rows+2b); src may alias its interior (the in-place call at level>0).
void copy_make_border_reflect101(const Img& src, Img& dst, int b) {
  const int w = src.cols, h = src.rows;
  for (int y = 0; y < h; y++) {  // interior rows first (safe when aliased: same bytes)
    const uint8_t* s = src.ptr(y);
    uint8_t*       d = dst.ptr(y + b);
    if (d + b != s) memmove(d + b, s, w);
    for (int x = 0; x < b; x++) d[x] = d[b + reflect101(x - b, w)];
    for (int x = 0; x < b; x++) d[b + w + x] = d[b + reflect101(w + x, w)];
  }
  for (int y = 0; y < b; y++) memcpy(dst.ptr(y), dst.ptr(b + reflect101(y - b, h)), w + 2 * b);
  for (int y = 0; y < b; y++)
    memcpy(dst.ptr(b + h + y), dst.ptr(b + reflect101(h + y, h)), w + 2 * b);
}

// Bresenham circle of radius 3, OpenCV order (features2d/fast_score.cpp, patternSize 16).
static const int kRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1},
                                 {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                 {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

// cornerScore<16>: the largest t for which p is still a FAST-9 corner = max(A,B) - 1,
//   A = max over the 16 arcs of 9 of min(v - ring), B = same with (ring - v).
int fast_corner_score(const uint8_t* p, size_t step) {
  int d[25];
  const int v = p[0];
  for (int k = 0; k < 25; k++)
    d[k] = v - p[(ptrdiff_t)kRing[k & 15][1] * (ptrdiff_t)step + kRing[k & 15][0]];
  int best = -1000;
  for (int k = 0; k < 16; k++) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; j++) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
    best = std::max(best, std::max(mn, -mx));
  }
  return best - 1;
}

// cv::FAST(roi, keypoints, threshold, nonmaxSuppression) with TYPE_9_16 (features2d/fast.cpp,
// FAST_t<16>), called at src/ORBextractor.cc:800,821.  Corners live on [3,cols-3)x[3,rows-3);
// stored score = cornerScore (uchar); NMS keeps strictly-greater-than-8-neighbours among the
// thresholded scores (0 elsewhere and outside the detection area); output row-major.
void fast9_16(const Img& im, std::vector<KeyPoint>& out, int threshold, bool nonmax) {
  out.clear();
  const int w = im.cols, h = im.rows;
  if (w < 7 || h < 7) return;
  std::vector<uint8_t> score((size_t)w * h, 0);
  std::vector<uint8_t> is_corner((size_t)w * h, 0);
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      const uint8_t* p = im.ptr(y) + x;
      const int v = p[0];
      {  // quick reject: an arc of 9 contains one pixel of every opposite pair
        const ptrdiff_t st = (ptrdiff_t)im.step;
        const int hi = v + threshold, lo = v - threshold;
        const int r0 = p[3 * st], r8 = p[-3 * st], r4 = p[3], r12 = p[-3];
        const bool b = (r0 > hi || r8 > hi) && (r4 > hi || r12 > hi);
        const bool d = (r0 < lo || r8 < lo) && (r4 < lo || r12 < lo);
        if (!b && !d) continue;
      }
      // 9 contiguous ring pixels all > v+t or all < v-t (strict)
      int run_b = 0, run_d = 0;
      bool corner = false;
      for (int k = 0; k < 25 && !corner; k++) {
        int r = p[(ptrdiff_t)kRing[k & 15][1] * (ptrdiff_t)im.step + kRing[k & 15][0]];
        run_b = (r > v + threshold) ? run_b + 1 : 0;
        run_d = (r < v - threshold) ? run_d + 1 : 0;
        corner = run_b >= 9 || run_d >= 9;
      }
      if (corner) {
        is_corner[(size_t)y * w + x] = 1;
        score[(size_t)y * w + x]     = (uint8_t)fast_corner_score(p, im.step);
      }
    }
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      if (!is_corner[(size_t)y * w + x]) continue;
      const int s = score[(size_t)y * w + x];
      bool keep = true;
      if (nonmax) {
        for (int dy = -1; dy <= 1 && keep; dy++)
          for (int dx = -1; dx <= 1; dx++)
            if ((dx || dy) && !(s > score[(size_t)(y + dy) * w + x + dx])) { keep = false; break; }
      }
      if (keep) {
        KeyPoint kp;
        kp.x = (float)x; kp.y = (float)y; kp.size = 7.f; kp.angle = -1.f; kp.response = (float)s;
        kp.octave = 0; kp.class_id = -1;
        out.push_back(kp);
      }
    }
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8U
// (src/ORBextractor.cc:1104).  OpenCV's 8-bit path is fixed point: the bit-exact kernel for
// n=7, sigma=2 in 8.8 format is {18,34,48,56,48,34,18} (sum 256); rows accumulate in 8.8
// (ufixedpoint16), columns in 16.16 (ufixedpoint32), final rounding (v + 32768) >> 16.
void gaussian_blur_7x7_s2(const Img& src, Img& dst) {
  static const int k[7] = {18, 34, 48, 56, 48, 34, 18};
  const int w = src.cols, h = src.rows;
  std::vector<uint16_t> tmp((size_t)w * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* s = src.ptr(y);
    for (int x = 0; x < w; x++) {
      int acc = 0;
      for (int i = 0; i < 7; i++) acc += k[i] * s[reflect101(x + i - 3, w)];
      tmp[(size_t)y * w + x] = (uint16_t)acc;
    }
  }
  for (int y = 0; y < h; y++) {
    uint8_t* d = dst.ptr(y);
    for (int x = 0; x < w; x++) {
      uint32_t acc = 0;
      for (int j = 0; j < 7; j++) acc += (uint32_t)k[j] * tmp[(size_t)reflect101(y + j - 3, h) * w + x];
      d[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

// scalar cv::fastAtan2 (core/mathfuncs_core.simd.hpp, atan_f32 reference form): degrees in
// [0,360), float32 Horner evaluation without FMA (this file is built -ffp-contract=off).
float fast_atan2(float y, float x) {
  static const float scale = (float)(180.0 / 3.14159265358979323846);
  static const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                     p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
  float ax = std::abs(x), ay = std::abs(y);
  float a, c, c2;
  if (ax >= ay) {
    c  = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a  = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c  = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a  = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ------------------------------------------------------------------------------------------
// ORBextractor (src/ORBextractor.cc)
// ------------------------------------------------------------------------------------------

static const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;  // :62-64

// IC_Angle (src/ORBextractor.cc:66-91)
static float IC_Angle(const Img& image, float ptx, float pty, const std::vector<int>& u_max) {
  int m_01 = 0, m_10 = 0;
  const uint8_t* center = image.ptr(cv_round(pty)) + cv_round(ptx);
  for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
  const ptrdiff_t step = (ptrdiff_t)image.step;
  for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
    int v_sum = 0;
    const int d = u_max[v];
    for (int u = -d; u <= d; ++u) {
      int val_plus = center[u + v * step], val_minus = center[u - v * step];
      v_sum += (val_plus - val_minus);
      m_10 += u * (val_plus + val_minus);
    }
    m_01 += v * v_sum;
  }
  return fast_atan2((float)m_01, (float)m_10);
}

// computeOrbDescriptor (src/ORBextractor.cc:93-139)
static const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
static void computeOrbDescriptor(const KeyPoint& kpt, const Img& img, uint8_t* desc) {
  const float angle = kpt.angle * factorPI;
  const float a = cosf(angle), b = sinf(angle);
  const uint8_t*  center = img.ptr(cv_round(kpt.y)) + cv_round(kpt.x);
  const ptrdiff_t step   = (ptrdiff_t)img.step;
  const signed char* pat = CUORB_BIT_PATTERN_31;
  for (int i = 0; i < 32; ++i) {
    int val = 0;
    for (int k = 0; k < 8; ++k, pat += 4) {
      const int x0 = pat[0], y0 = pat[1], x1 = pat[2], y1 = pat[3];
      const int t0 = center[cv_round(x0 * b + y0 * a) * step + cv_round(x0 * a - y0 * b)];
      const int t1 = center[cv_round(x1 * b + y1 * a) * step + cv_round(x1 * a - y1 * b)];
      val |= (t0 < t1) << k;
    }
    desc[i] = (uint8_t)val;
  }
}

// ctor (src/ORBextractor.cc:403-465)
ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST,
                           int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST),
      minThFAST(_minThFAST) {
  mvScaleFactor.resize(nlevels);
  mvLevelSigma2.resize(nlevels);
  mvScaleFactor[0] = 1.0f;
  mvLevelSigma2[0] = 1.0f;
  for (int i = 1; i < nlevels; i++) {
    mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;  // float*double -> float
    mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
  }
  mvInvScaleFactor.resize(nlevels);
  mvInvLevelSigma2.resize(nlevels);
  for (int i = 0; i < nlevels; i++) {
    mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
    mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
  }
  mnFeaturesPerLevel.resize(nlevels);
  float factor = 1.0f / scaleFactor;
  float nDesiredFeaturesPerScale =
      nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
  int sumFeatures = 0;
  for (int level = 0; level < nlevels - 1; level++) {
    mnFeaturesPerLevel[level] = cv_round(nDesiredFeaturesPerScale);
    sumFeatures += mnFeaturesPerLevel[level];
    nDesiredFeaturesPerScale *= factor;
  }
  mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);

  umax.resize(HALF_PATCH_SIZE + 1);
  int v, v0, vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
  int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
  const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
  for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
  for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
}

// ComputePyramid (src/ORBextractor.cc:1140-1173)
void ORBextractor::ComputePyramid(const Img& image) {
  pyr.storage.resize(nlevels);
  pyr.level.resize(nlevels);
  pyr.whole.resize(nlevels);
  for (int level = 0; level < nlevels; ++level) {
    float scale = mvInvScaleFactor[level];
    int   sw = cv_round((float)image.cols * scale), sh = cv_round((float)image.rows * scale);
    int   ww = sw + EDGE_THRESHOLD * 2, wh = sh + EDGE_THRESHOLD * 2;
    pyr.storage[level].assign((size_t)ww * wh, 0);
    Img temp;
    temp.data = pyr.storage[level].data(); temp.cols = ww; temp.rows = wh; temp.step = ww;
    Img roi;
    roi.data = temp.data + (size_t)EDGE_THRESHOLD * ww + EDGE_THRESHOLD;
    roi.cols = sw; roi.rows = sh; roi.step = ww;
    pyr.whole[level] = temp;
    pyr.level[level] = roi;
    if (level != 0) {
      resize_linear_8u(pyr.level[level - 1], pyr.level[level]);
      copy_make_border_reflect101(pyr.level[level], temp, EDGE_THRESHOLD);
    } else {
      copy_make_border_reflect101(image, temp, EDGE_THRESHOLD);
    }
  }
}

// ExtractorNode (include/ORBextractor.h:29-39) and DivideNode (src/ORBextractor.cc:481-540)
namespace {
struct P2i { int x = 0, y = 0; };
struct ExtractorNode {
  std::vector<KeyPoint> vKeys;
  P2i UL, UR, BL, BR;
  std::list<ExtractorNode>::iterator lit;
  bool bNoMore = false;
  void DivideNode(ExtractorNode& n1, ExtractorNode& n2, ExtractorNode& n3, ExtractorNode& n4);
};

void ExtractorNode::DivideNode(ExtractorNode& n1, ExtractorNode& n2, ExtractorNode& n3,
                               ExtractorNode& n4) {
  const int halfX = (int)std::ceil(static_cast<float>(UR.x - UL.x) / 2);
  const int halfY = (int)std::ceil(static_cast<float>(BR.y - UL.y) / 2);
  n1.UL = UL;                       n1.UR = {UL.x + halfX, UL.y};
  n1.BL = {UL.x, UL.y + halfY};     n1.BR = {UL.x + halfX, UL.y + halfY};
  n2.UL = n1.UR; n2.UR = UR; n2.BL = n1.BR; n2.BR = {UR.x, UL.y + halfY};
  n3.UL = n1.BL; n3.UR = n1.BR; n3.BL = BL; n3.BR = {n1.BR.x, BL.y};
  n4.UL = n3.UR; n4.UR = n2.BR; n4.BL = n3.BR; n4.BR = BR;
  for (size_t i = 0; i < vKeys.size(); i++) {
    const KeyPoint& kp = vKeys[i];
    if (kp.x < n1.UR.x) {
      if (kp.y < n1.BR.y) n1.vKeys.push_back(kp);
      else n3.vKeys.push_back(kp);
    } else if (kp.y < n1.BR.y) n2.vKeys.push_back(kp);
    else n4.vKeys.push_back(kp);
  }
  if (n1.vKeys.size() == 1) n1.bNoMore = true;
  if (n2.vKeys.size() == 1) n2.bNoMore = true;
  if (n3.vKeys.size() == 1) n3.bNoMore = true;
  if (n4.vKeys.size() == 1) n4.bNoMore = true;
}

// compareNodes (src/ORBextractor.cc:542-554): strict weak order on (nKeys, UL.x); ties exist, and
// std::sort is not stable, so libstdc++'s introsort placement is part of the behaviour.
bool compareNodes(std::pair<int, ExtractorNode*>& e1, std::pair<int, ExtractorNode*>& e2) {
  if (e1.first < e2.first) return true;
  if (e1.first > e2.first) return false;
  return e1.second->UL.x < e2.second->UL.x;
}
}  // namespace

// DistributeOctTree (src/ORBextractor.cc:556-753)
std::vector<KeyPoint> ORBextractor::DistributeOctTree(const std::vector<KeyPoint>& vToDistributeKeys,
                                                      int minX, int maxX, int minY, int maxY,
                                                      int N, int /*level*/) {
  const int   nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
  const float hX   = static_cast<float>(maxX - minX) / nIni;

  std::list<ExtractorNode>    lNodes;
  std::vector<ExtractorNode*> vpIniNodes(nIni);
  for (int i = 0; i < nIni; i++) {
    ExtractorNode ni;
    ni.UL = {(int)(hX * static_cast<float>(i)), 0};
    ni.UR = {(int)(hX * static_cast<float>(i + 1)), 0};
    ni.BL = {ni.UL.x, maxY - minY};
    ni.BR = {ni.UR.x, maxY - minY};
    lNodes.push_back(ni);
    vpIniNodes[i] = &lNodes.back();
  }
  for (size_t i = 0; i < vToDistributeKeys.size(); i++) {
    const KeyPoint& kp = vToDistributeKeys[i];
    vpIniNodes[(int)(kp.x / hX)]->vKeys.push_back(kp);
  }
  auto lit = lNodes.begin();
  while (lit != lNodes.end()) {
    if (lit->vKeys.size() == 1) { lit->bNoMore = true; lit++; }
    else if (lit->vKeys.empty()) lit = lNodes.erase(lit);
    else lit++;
  }

  bool bFinish = false;
  std::vector<std::pair<int, ExtractorNode*>> vSizeAndPointerToNode;
  auto add_children = [&](ExtractorNode* n[4], int* nToExpand) {
    for (int c = 0; c < 4; c++) {
      if (n[c]->vKeys.size() > 0) {
        lNodes.push_front(*n[c]);
        if (n[c]->vKeys.size() > 1) {
          if (nToExpand) (*nToExpand)++;
          vSizeAndPointerToNode.push_back(std::make_pair((int)n[c]->vKeys.size(), &lNodes.front()));
          lNodes.front().lit = lNodes.begin();
        }
      }
    }
  };

  while (!bFinish) {
    int prevSize = (int)lNodes.size();
    lit = lNodes.begin();
    int nToExpand = 0;
    vSizeAndPointerToNode.clear();
    while (lit != lNodes.end()) {
      if (lit->bNoMore) { lit++; continue; }
      ExtractorNode n1, n2, n3, n4;
      lit->DivideNode(n1, n2, n3, n4);
      ExtractorNode* ch[4] = {&n1, &n2, &n3, &n4};
      add_children(ch, &nToExpand);
      lit = lNodes.erase(lit);
    }
    if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
      bFinish = true;
    } else if (((int)lNodes.size() + nToExpand * 3) > N) {
      while (!bFinish) {
        prevSize = (int)lNodes.size();
        std::vector<std::pair<int, ExtractorNode*>> vPrev = vSizeAndPointerToNode;
        vSizeAndPointerToNode.clear();
        std::sort(vPrev.begin(), vPrev.end(), compareNodes);
        for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
          ExtractorNode n1, n2, n3, n4;
          vPrev[j].second->DivideNode(n1, n2, n3, n4);
          ExtractorNode* ch[4] = {&n1, &n2, &n3, &n4};
          add_children(ch, nullptr);
          lNodes.erase(vPrev[j].second->lit);
          if ((int)lNodes.size() >= N) break;
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
      }
    }
  }

  std::vector<KeyPoint> vResultKeys;
  vResultKeys.reserve(nfeatures);
  for (auto it = lNodes.begin(); it != lNodes.end(); it++) {
    std::vector<KeyPoint>& vNodeKeys = it->vKeys;
    KeyPoint* pKP         = &vNodeKeys[0];
    float     maxResponse = pKP->response;
    for (size_t k = 1; k < vNodeKeys.size(); k++)
      if (vNodeKeys[k].response > maxResponse) { pKP = &vNodeKeys[k]; maxResponse = vNodeKeys[k].response; }
    vResultKeys.push_back(*pKP);
  }
  return vResultKeys;
}

// ComputeKeyPointsOctTree (src/ORBextractor.cc:755-881)
void ORBextractor::ComputeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& allKeypoints) {
  allKeypoints.resize(nlevels);
  dbg_candidates.assign(nlevels, {});
  const float W = 35;
  for (int level = 0; level < nlevels; ++level) {
    const int minBorderX = EDGE_THRESHOLD - 3;
    const int minBorderY = minBorderX;
    const int maxBorderX = pyr.level[level].cols - EDGE_THRESHOLD + 3;
    const int maxBorderY = pyr.level[level].rows - EDGE_THRESHOLD + 3;

    std::vector<KeyPoint> vToDistributeKeys;
    const float width  = (float)(maxBorderX - minBorderX);
    const float height = (float)(maxBorderY - minBorderY);
    const int nCols = (int)(width / W);
    const int nRows = (int)(height / W);
    const int wCell = (int)std::ceil(width / nCols);
    const int hCell = (int)std::ceil(height / nRows);

    for (int i = 0; i < nRows; i++) {
      const float iniY = (float)(minBorderY + i * hCell);
      float       maxY = iniY + hCell + 6;
      if (iniY >= maxBorderY - 3) continue;
      if (maxY > maxBorderY) maxY = (float)maxBorderY;
      for (int j = 0; j < nCols; j++) {
        const float iniX = (float)(minBorderX + j * wCell);
        float       maxX = iniX + wCell + 6;
        if (iniX >= maxBorderX - 6) continue;
        if (maxX > maxBorderX) maxX = (float)maxBorderX;

        Img roi;
        roi.data = pyr.level[level].ptr((int)iniY) + (int)iniX;
        roi.cols = (int)maxX - (int)iniX;
        roi.rows = (int)maxY - (int)iniY;
        roi.step = pyr.level[level].step;
        std::vector<KeyPoint> vKeysCell;
        fast9_16(roi, vKeysCell, iniThFAST, true);
        if (vKeysCell.empty()) fast9_16(roi, vKeysCell, minThFAST, true);
        for (auto& kp : vKeysCell) {
          kp.x += j * wCell;
          kp.y += i * hCell;
          vToDistributeKeys.push_back(kp);
        }
      }
    }
    dbg_candidates[level] = vToDistributeKeys;

    std::vector<KeyPoint>& keypoints = allKeypoints[level];
    keypoints = DistributeOctTree(vToDistributeKeys, minBorderX, maxBorderX, minBorderY,
                                  maxBorderY, mnFeaturesPerLevel[level], level);
    const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
    for (auto& kp : keypoints) {
      kp.x += minBorderX;
      kp.y += minBorderY;
      kp.octave = level;
      kp.size   = (float)scaledPatchSize;
    }
  }
  for (int level = 0; level < nlevels; ++level)  // computeOrientation (:470-479)
    for (auto& kp : allKeypoints[level]) kp.angle = IC_Angle(pyr.level[level], kp.x, kp.y, umax);
  dbg_level_keys = allKeypoints;  // level coordinates, before operator() scales pt
}

// operator() (src/ORBextractor.cc:1052-1138).  Returns monoIndex, or -1 for an empty image.
int ORBextractor::operator()(const Img& image, std::vector<KeyPoint>& _keypoints,
                             std::vector<uint8_t>& _descriptors, const int vLappingArea[2]) {
  if (image.data == nullptr || image.cols == 0 || image.rows == 0) return -1;
  ComputePyramid(image);
  std::vector<std::vector<KeyPoint>> allKeypoints;
  ComputeKeyPointsOctTree(allKeypoints);

  int nkeypoints = 0;
  for (int level = 0; level < nlevels; ++level) nkeypoints += (int)allKeypoints[level].size();
  _descriptors.assign((size_t)nkeypoints * 32, 0);
  _keypoints.assign(nkeypoints, KeyPoint());
  dbg_blurred.assign(nlevels, {});

  int monoIndex = 0, stereoIndex = nkeypoints - 1;
  for (int level = 0; level < nlevels; ++level) {
    std::vector<KeyPoint>& keypoints = allKeypoints[level];
    const int nkeypointsLevel = (int)keypoints.size();
    if (nkeypointsLevel == 0) continue;

    const Img& lv = pyr.level[level];
    dbg_blurred[level].assign((size_t)lv.cols * lv.rows, 0);
    Img working;
    working.data = dbg_blurred[level].data(); working.cols = lv.cols; working.rows = lv.rows;
    working.step = lv.cols;
    gaussian_blur_7x7_s2(lv, working);  // clone() + GaussianBlur(7x7, 2, 2, REFLECT_101)

    std::vector<uint8_t> desc((size_t)nkeypointsLevel * 32);
    for (int i = 0; i < nkeypointsLevel; i++) computeOrbDescriptor(keypoints[i], working, &desc[(size_t)i * 32]);

    const float scale = mvScaleFactor[level];
    int i = 0;
    for (auto& kp : keypoints) {
      if (level != 0) { kp.x *= scale; kp.y *= scale; }
      if (kp.x >= vLappingArea[0] && kp.x <= vLappingArea[1]) {
        _keypoints.at(stereoIndex) = kp;
        memcpy(&_descriptors[(size_t)stereoIndex * 32], &desc[(size_t)i * 32], 32);
        stereoIndex--;
      } else {
        _keypoints.at(monoIndex) = kp;
        memcpy(&_descriptors[(size_t)monoIndex * 32], &desc[(size_t)i * 32], 32);
        monoIndex++;
      }
      i++;
    }
  }
  return monoIndex;
}

}  // namespace orb_oracle
