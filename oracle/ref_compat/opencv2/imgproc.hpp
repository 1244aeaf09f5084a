// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stand-in for <opencv2/imgproc.hpp>; see core.hpp.
// cv::resize (src/ORBextractor.cc:1150) and cv::GaussianBlur (:1104) forward to the restated,
// cv2-pinned primitives of oracle/orb_oracle.cpp; any use outside what those cover aborts loudly.
#pragma once
#include "core.hpp"

namespace cv {
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };

void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0,
            int interpolation = INTER_LINEAR);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);
}  // namespace cv
