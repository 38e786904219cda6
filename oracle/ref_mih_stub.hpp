// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// Force-included (-include) in front of /root/reference/src/Hashing.cc, which is compiled UNMODIFIED: defining the
// include guard of the reference's include/MapPoint.h keeps that header (and through it KeyFrame.h, Frame.h, DBoW2,
// Eigen, g2o -- none of which this image has) out, and the stand-in below declares the four MapPoint members
// HASHING::MultiIndexHashing touches (src/Hashing.cc:49, :541-542, :571-589) with the reference's names and types
// (include/MapPoint.h:96, :152-159; MapPoint::GetDescriptor returns a clone, src/MapPoint.cc:460-464).
#pragma once
#define MAPPOINT_H
#include <cmath>
#include <climits>
#include <cstdlib>
#include <iostream>
#include <numeric>
#include <opencv2/core/core.hpp>      // oracle/ref_shim
#include <vector>

using namespace std;                  // the reference's headers do this long before Hashing.cc is reached

namespace ORB_SLAM2 {
class MapPoint {
public:
    long unsigned int mnId = 0;
    std::vector<bool> mvbHashed = std::vector<bool>(32, false);      // NUM_TOTAL_HASHTABLES, src/MapPoint.cc:54
    int mnQueriedScore = 0;
    cv::Mat GetDescriptor() { return desc_.clone(); }
    cv::Mat desc_;
};
}  // namespace ORB_SLAM2
