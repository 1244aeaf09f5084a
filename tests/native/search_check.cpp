// Exercises the shim's non-tracking ORBmatcher searches (shim/ORBmatcher.h) on small Frame / KeyFrame
// structs that carry the reference's member names: the monocular-initialisation flow
// (Tracking::MonocularInitialization, src/Tracking.cc:2228-2262: SearchForInitialization on two frames, twice),
// a Sim3-style SearchByProjection into a KeyFrame, Fuse, and SearchForTriangulation.  Results go to a
// binary file that the pytest compares with the CPU oracle fed with the same rule-built inputs.
// usage: search_check a.raw b.raw w h nfeatures out.bin
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

#include "../../shim/ORBextractor.h"
#include "../../shim/ORBmatcher.h"

struct MiniPoint { int id; };

struct MiniKF {  // members of Frame / KeyFrame the searches read (include/Frame.h, include/KeyFrame.h)
  int N = 0, Nleft = -1, NLeft = -1;
  std::vector<cv::KeyPoint> mvKeysUn, mvKeys, mvKeysRight;   // mvKeys / mvKeysRight: only read for rig frames (NLeft != -1)
  cv::Mat mDescriptors;
  std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0, mfGridElementWidthInv = 0, mfGridElementHeightInv = 0, mb = 0.1f, mbf = 40.f;
  std::vector<MiniPoint*> mvpMapPoints;
  std::map<unsigned, std::vector<unsigned>> mFeatVec;
  MiniPoint* GetMapPoint(int i) const { return mvpMapPoints[i]; }
  std::vector<MiniPoint*> GetMapPointMatches() const { return mvpMapPoints; }
};

static cv::Mat load(const char* path, int w, int h) {
  cv::Mat m(h, w, CV_8UC1);
  FILE* f = fopen(path, "rb");
  if (!f || fread(m.data, 1, (size_t)w * h, f) != (size_t)w * h) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
  fclose(f);
  return m;
}

static void fill(MiniKF& F, ORB_SLAM3::ORBextractor& ex, const cv::Mat& im, int w, int h) {
  std::vector<int> lap = {0, 0};
  ex(im, cv::Mat(), F.mvKeysUn, F.mDescriptors, lap);
  F.N = (int)F.mvKeysUn.size();
  F.mvScaleFactors = ex.GetScaleFactors();
  F.mvLevelSigma2 = ex.GetScaleSigmaSquares();
  F.mvInvLevelSigma2 = ex.GetInverseScaleSigmaSquares();
  F.mnMaxX = (float)w; F.mnMaxY = (float)h;
  F.mfGridElementWidthInv = 64.0f / (F.mnMaxX - F.mnMinX);      // src/Frame.cc:240-243
  F.mfGridElementHeightInv = 48.0f / (F.mnMaxY - F.mnMinY);
  F.mvpMapPoints.assign(F.N, nullptr);
  for (int i = 0; i < F.N; i++) F.mFeatVec[(unsigned)F.mvKeysUn[i].pt.y / 10u].push_back((unsigned)i);
}

int main(int argc, char** argv) {
  if (argc != 7) { fprintf(stderr, "usage\n"); return 2; }
  const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
  cv::Mat imA = load(argv[1], w, h), imB = load(argv[2], w, h);
  ORB_SLAM3::ORBextractor exA(nf, 1.2f, 8, 20, 7, w, h), exB(nf, 1.2f, 8, 20, 7, w, h);
  MiniKF A, B;
  fill(A, exA, imA, w, h);
  fill(B, exB, imB, w, h);
  ORB_SLAM3::ORBmatcher matcher(0.9f, true);

  // monocular initialisation: vbPrevMatched starts at the first frame's keypoints (src/Tracking.cc:2236-2239)
  std::vector<cv::Point2f> prev(A.N);
  for (int i = 0; i < A.N; i++) prev[i] = A.mvKeysUn[i].pt;
  std::vector<int> ini1, ini2;
  const int n_ini1 = matcher.SearchForInitialization(A, B, prev, ini1, 100);
  const int n_ini2 = matcher.SearchForInitialization(A, B, prev, ini2, 100);

  // every feature of A becomes a MapPoint projected near itself (rule mirrored in the pytest)
  std::vector<MiniPoint> pts(A.N);
  std::vector<MiniPoint*> vpPoints(A.N);
  ORB_SLAM3::ORBmatcher::ProjectedPoints P;
  for (int i = 0; i < A.N; i++) {
    pts[i].id = i; vpPoints[i] = &pts[i];
    const cv::KeyPoint& kp = A.mvKeysUn[i];
    const int lvl = kp.octave + (i & 1) > 7 ? 7 : kp.octave + (i & 1);
    P.push(i % 7 != 0, kp.pt.x + 1.5f, kp.pt.y - 0.75f, lvl, A.mDescriptors.ptr<unsigned char>(i), kp.pt.x - 9.0f, kp.angle);
  }
  std::vector<MiniPoint*> vpMatched(A.N, nullptr);
  for (int i = 0; i < A.N; i += 5) vpMatched[i] = &pts[0];       // already matched on entry
  std::vector<int> sim3(A.N, -1);
  const int n_sim3 = matcher.SearchByProjection(&A, P, vpPoints, vpMatched, 10, 1.0f);
  for (int i = 0; i < A.N; i++)
    if (vpMatched[i] && !(i % 5 == 0 && vpMatched[i] == &pts[0])) sim3[i] = vpMatched[i]->id;
  std::vector<int> fused;
  const int n_fuse = matcher.Fuse(&A, P, 3.0f, true, fused);

  // triangulation between the two views: rectified-pair geometry
  const float F12[9] = {0.f, 1e-4f, 0.f, -2e-4f, 0.f, -1.f, 3e-4f, 1.f, 0.f};
  const float ep[2] = {(float)w * 0.5f, (float)h * 0.5f};
  for (int i = 0; i < A.N; i += 3) A.mvpMapPoints[i] = &pts[0];
  std::vector<std::pair<std::size_t, std::size_t>> pairs;
  const int n_tri = matcher.SearchForTriangulation(&A, &B, F12, ep, pairs, false, false);
  std::vector<int> tri(A.N, -1);
  for (auto& pr : pairs) tri[pr.first] = (int)pr.second;

  FILE* f = fopen(argv[6], "wb");
  int hdr[8] = {A.N, B.N, n_ini1, n_ini2, n_sim3, n_fuse, n_tri, 0};
  fwrite(hdr, sizeof(int), 8, f);
  fwrite(A.mvKeysUn.data(), sizeof(cv::KeyPoint), A.N, f);
  fwrite(A.mDescriptors.data, 32, A.N, f);
  fwrite(B.mvKeysUn.data(), sizeof(cv::KeyPoint), B.N, f);
  fwrite(B.mDescriptors.data, 32, B.N, f);
  fwrite(ini1.data(), 4, A.N, f);
  fwrite(ini2.data(), 4, A.N, f);
  fwrite(prev.data(), 8, A.N, f);
  fwrite(sim3.data(), 4, A.N, f);
  fwrite(fused.data(), 4, A.N, f);
  fwrite(tri.data(), 4, A.N, f);
  fclose(f);
  printf("search shim ok: %d/%d init, %d sim3, %d fuse, %d triangulation\n", n_ini1, n_ini2, n_sim3, n_fuse, n_tri);
  return 0;
}
