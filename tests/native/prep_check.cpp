// Exercises shim/FramePrep.h the way System::TrackStereo + Tracking::GrabImageStereo prepare a frame
// (src/System.cc:280-300, src/Tracking.cc:1339-1357): rectify a BGR frame with CV_32FC1 maps, convert
// to gray, and (fused form) extract.  usage: prep_check frame.raw w h mapx.raw mapy.raw nfeatures out.bin
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../shim/FramePrep.h"
#include "../../shim/ORBextractor.h"

static void slurp(const char* path, void* dst, size_t n) {
  FILE* f = fopen(path, "rb");
  if (!f || fread(dst, 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc != 8) { fprintf(stderr, "usage\n"); return 2; }
  const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[6]);
  cv::Mat im(h, w, CV_8UC3), M1(h, w, CV_32FC1), M2(h, w, CV_32FC1);
  slurp(argv[1], im.data, (size_t)w * h * 3);
  slurp(argv[4], M1.data, (size_t)w * h * 4);
  slurp(argv[5], M2.data, (size_t)w * h * 4);
  ORB_SLAM3::FramePrep prep(w, h, M1, M2);
  ORB_SLAM3::ORBextractor ex(nf, 1.2f, 8, 20, 7, w, h);
  cv::Mat gray;
  prep(im, /*bRGB=*/false, gray);
  std::vector<cv::KeyPoint> k0, k1;
  cv::Mat d0, d1;
  std::vector<int> lap = {0, 0};
  const int mono0 = ex(gray, cv::Mat(), k0, d0, lap);                  // two-step, as the reference does it
  const int mono1 = prep.Extract(ex, im, false, k1, d1, lap);          // fused
  if (mono0 != mono1 || k0.size() != k1.size()) return 3;
  for (size_t i = 0; i < k0.size(); i++)
    if (memcmp(&k0[i], &k1[i], sizeof(cv::KeyPoint)) != 0 || memcmp(d0.ptr<unsigned char>((int)i), d1.ptr<unsigned char>((int)i), 32) != 0) return 4;
  FILE* f = fopen(argv[7], "wb");
  int hdr[2] = {(int)k1.size(), mono1};
  fwrite(hdr, sizeof(int), 2, f);
  fwrite(gray.data, 1, (size_t)w * h, f);
  fwrite(k1.data(), sizeof(cv::KeyPoint), k1.size(), f);
  fwrite(d1.data, 32, k1.size(), f);
  fclose(f);
  printf("prep ok: %d keypoints\n", hdr[0]);
  return 0;
}
