// Exercises the C++ drop-in shim (shim/ORBextractor.h, shim/ORBmatcher.h) the way the reference's
// Frame constructor does (src/Frame.cc:202-225): left/right operator(), ComputeStereoMatches,
// and DescriptorDistance.  Writes the results to a binary file the pytest compares with the
// oracle.  usage: shim_check left.raw right.raw w h nfeatures mb mbf out.bin
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../shim/ORBextractor.h"
#include "../../shim/ORBmatcher.h"

struct MiniFrame {  // the Frame members ComputeStereoMatches touches (include/Frame.h)
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
  cv::Mat mDescriptors, mDescriptorsRight;
  std::vector<float> mvuRight, mvDepth;
  float mb, mbf;
};

static cv::Mat load(const char* path, int w, int h) {
  cv::Mat m(h, w, CV_8UC1);
  FILE* f = fopen(path, "rb");
  if (!f || fread(m.data, 1, (size_t)w * h, f) != (size_t)w * h) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
  fclose(f);
  return m;
}

int main(int argc, char** argv) {
  if (argc != 9) { fprintf(stderr, "usage\n"); return 2; }
  const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
  cv::Mat imL = load(argv[1], w, h), imR = load(argv[2], w, h);
  ORB_SLAM3::ORBextractor exL(nf, 1.2f, 8, 20, 7, w, h), exR(nf, 1.2f, 8, 20, 7, w, h);
  MiniFrame F;
  F.mb = (float)atof(argv[6]); F.mbf = (float)atof(argv[7]);
  std::vector<int> lap = {0, 0};
  const int monoL = exL(imL, cv::Mat(), F.mvKeys, F.mDescriptors, lap);
  const int monoR = exR(imR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight, lap);
  // empty image -> -1, outputs untouched (src/ORBextractor.cc:1060-1062)
  std::vector<cv::KeyPoint> untouched(3);
  cv::Mat dummy;
  if (exL(cv::Mat(), cv::Mat(), untouched, dummy, lap) != -1 || untouched.size() != 3) return 3;
  // re-run the left image: the empty call must not have disturbed the handle
  std::vector<cv::KeyPoint> again; cv::Mat dagain;
  if (exL(imL, cv::Mat(), again, dagain, lap) != monoL || again.size() != F.mvKeys.size()) return 4;
  ORB_SLAM3::ORBmatcher::ComputeStereoMatches(F, exL, exR);
  {  // the fused call leaves exactly what the three steps left
    MiniFrame G;
    G.mb = F.mb; G.mbf = F.mbf;
    const int monoG = ORB_SLAM3::ORBmatcher::ExtractStereo(G, exL, exR, imL, imR);
    bool same = monoG == monoL && G.mvKeys.size() == F.mvKeys.size() && G.mvKeysRight.size() == F.mvKeysRight.size() &&
                G.mvuRight.size() == F.mvuRight.size();
    same = same && memcmp(G.mvKeys.data(), F.mvKeys.data(), sizeof(cv::KeyPoint) * F.mvKeys.size()) == 0;
    same = same && memcmp(G.mvKeysRight.data(), F.mvKeysRight.data(), sizeof(cv::KeyPoint) * F.mvKeysRight.size()) == 0;
    for (size_t i = 0; same && i < F.mvKeys.size(); i++) same = memcmp(G.mDescriptors.ptr<unsigned char>((int)i), F.mDescriptors.ptr<unsigned char>((int)i), 32) == 0;
    for (size_t i = 0; same && i < F.mvKeysRight.size(); i++) same = memcmp(G.mDescriptorsRight.ptr<unsigned char>((int)i), F.mDescriptorsRight.ptr<unsigned char>((int)i), 32) == 0;
    same = same && memcmp(G.mvuRight.data(), F.mvuRight.data(), 4 * F.mvuRight.size()) == 0 &&
           memcmp(G.mvDepth.data(), F.mvDepth.data(), 4 * F.mvDepth.size()) == 0;
    if (!same) { fprintf(stderr, "ExtractStereo differs from ExtractORB x2 + ComputeStereoMatches\n"); return 6; }
  }
  exL.SyncImagePyramid();
  int selfdist = ORB_SLAM3::ORBmatcher::DescriptorDistance(F.mDescriptors.row(0), F.mDescriptors.row(0));
  int d01 = F.mvKeys.size() > 1 ? ORB_SLAM3::ORBmatcher::DescriptorDistance(F.mDescriptors.row(0), F.mDescriptors.row(1)) : 0;
  FILE* f = fopen(argv[8], "wb");
  int hdr[8] = {(int)F.mvKeys.size(), (int)F.mvKeysRight.size(), monoL, monoR, selfdist, d01,
                exL.mvImagePyramid[1].cols, exL.mvImagePyramid[1].rows};
  fwrite(hdr, sizeof(int), 8, f);
  fwrite(F.mvKeys.data(), sizeof(cv::KeyPoint), F.mvKeys.size(), f);
  fwrite(F.mDescriptors.data, 32, F.mvKeys.size(), f);
  fwrite(F.mvKeysRight.data(), sizeof(cv::KeyPoint), F.mvKeysRight.size(), f);
  fwrite(F.mDescriptorsRight.data, 32, F.mvKeysRight.size(), f);
  fwrite(F.mvuRight.data(), 4, F.mvuRight.size(), f);
  fwrite(F.mvDepth.data(), 4, F.mvDepth.size(), f);
  fwrite(exL.mvImagePyramid[1].data, 1, (size_t)hdr[6] * hdr[7], f);
  fclose(f);
  printf("shim ok: %d + %d keypoints\n", hdr[0], hdr[1]);
  return 0;
}
