// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Translation unit that gives the stub classes of oracle/ref_compat the reference's OWN function
// bodies: oracle/ref_extract.py lifts them verbatim from /root/reference/src at build time into
// oracle/_ref/gen/*.inc (git-ignored, `#line`-annotated) and they are compiled here unchanged.
#include <algorithm>
#include <climits>
#include <cmath>
#include <mutex>
#include <thread>
#include <vector>

#include "Frame.h"
#include "GeometricCamera.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "ORBextractor.h"  // the reference's (-I $(REF)/include)
#include "ORBmatcher.h"    // the reference's

namespace ORB_SLAM3 {

// static data members of Frame (src/Frame.cc:43-50)
float Frame::fx = 0, Frame::fy = 0, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 0, Frame::invfy = 0;
float Frame::mnMinX = 0, Frame::mnMinY = 0, Frame::mnMaxX = 0, Frame::mnMaxY = 0;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
cv::BFMatcher Frame::BFmatcher = cv::BFMatcher(cv::NORM_HAMMING);   // src/Frame.cc:53

#include "gen/frame_lifted.inc"
#include "gen/keyframe_lifted.inc"
#include "gen/mappoint_lifted.inc"
#include "gen/pinhole_lifted.inc"

}  // namespace ORB_SLAM3
