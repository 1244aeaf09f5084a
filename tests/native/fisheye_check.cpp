// Drives shim/FrameOps.h's ComputeStereoFishEyeMatches<KannalaBrandt8>(Frame&) over the stand-in Frame / camera classes
// of oracle/ref_compat (headers only; nothing of the reference is linked) with the deterministic triangulation of
// fake_triangulate.h, and writes mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints to a file the pytest
// compares with the reference's own function (src/Frame.cc:1303-1355, lifted into oracle/_ref).
// usage: fisheye_check case.bin out.bin
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "Frame.h"
#include "GeometricCamera.h"
#include "fake_triangulate.h"

#include "../../shim/FrameOps.h"

namespace ORB_SLAM3 {
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMinY = 0, Frame::mnMaxX = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
}  // namespace ORB_SLAM3
using namespace ORB_SLAM3;

template <class T> static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
  return v;
}

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: fisheye_check case.bin out.bin\n"); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  const std::vector<int32_t> hi = rd<int32_t>(f, 5);      // nl, nr, monoLeft, monoRight, nlevels
  const int nl = hi[0], nr = hi[1];
  Frame F;
  KannalaBrandt8 cam, cam2;
  cam.triangulate_hook = fake_triangulate;
  F.mpCamera = &cam; F.mpCamera2 = &cam2;
  F.Nleft = nl; F.Nright = nr; F.N = nl + nr; F.monoLeft = hi[2]; F.monoRight = hi[3];
  F.mvKeys = rd<cv::KeyPoint>(f, nl); F.mvKeysRight = rd<cv::KeyPoint>(f, nr);
  std::vector<uint8_t> dl = rd<uint8_t>(f, (size_t)nl * 32), dr = rd<uint8_t>(f, (size_t)nr * 32);
  F.mDescriptors = cv::Mat(nl, 32, CV_8UC1, dl.data(), 32);
  F.mDescriptorsRight = cv::Mat(nr, 32, CV_8UC1, dr.data(), 32);
  F.mvLevelSigma2 = rd<float>(f, hi[4]);
  fclose(f);
  ComputeStereoFishEyeMatches<KannalaBrandt8>(F);
  f = fopen(argv[2], "wb");
  fwrite(F.mvLeftToRightMatch.data(), 4, nl, f);
  fwrite(F.mvRightToLeftMatch.data(), 4, nr, f);
  fwrite(F.mvDepth.data(), 4, nl, f);
  for (int i = 0; i < nl; i++) { float p[3] = {F.mvStereo3Dpoints[i](0), F.mvStereo3Dpoints[i](1), F.mvStereo3Dpoints[i](2)}; fwrite(p, 4, 3, f); }
  fclose(f);
  int n = 0;
  for (int v : F.mvLeftToRightMatch) n += v >= 0;
  printf("fisheye stereo ok: %d matches\n", n);
  return 0;
}
