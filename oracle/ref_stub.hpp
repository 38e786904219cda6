// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// Stand-in declarations of ORB_SLAM2::MapPoint, ::Frame and ::ORBmatcher, holding exactly the members that the
// reference function bodies lifted by oracle/ref_extract.py read or write -- same names, same types, same
// static-ness as /root/reference/include/Frame.h:283-393, include/ORBmatcher.h:44-308, include/MapPoint.h.  The
// real headers cannot be included here: they pull in DBoW2's vocabulary, Eigen, Armadillo, g2o and ROS.
// What is DECLARED here is the builder's; every function BODY compiled against it is the reference's own text
// (oracle/_ref/gen/*.inc, produced at build time, never committed).
#pragma once
#include <climits>
#include <cmath>
#include <list>
#include <map>
#include <utility>
#include <vector>

#include "ORBextractor.h"   // the reference's own header (-I/root/reference/include)

// include/Frame.h:46, :92-93
#define DISPARITY_THRES 50.0
#define FRAME_GRID_COLS 64
#define FRAME_GRID_ROWS 48

namespace ORB_SLAM2 {

// include/MapPoint.h: only the tracking fields and accessors the lifted matcher / stereo bodies touch
class MapPoint {
public:
    bool mbTrackInView = true;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 1.f, mTrackProjX = 0.f, mTrackProjY = 0.f, mTrackProjXR = 0.f;

    bool isBad() { return bad_; }
    int Observations() { return nobs_; }
    cv::Mat GetDescriptor() { return desc_.clone(); }
    cv::Mat GetWorldPos() { return world_.clone(); }

    bool bad_ = false;
    int nobs_ = 0;
    cv::Mat desc_, world_;
    int tag_ = -1;          // index of this point in the caller's array (to report which point took a key)
};

class Frame {
public:
    Frame() : mpORBextractorLeft(nullptr), mpORBextractorRight(nullptr), mbf(0), mb(0), N(0) {}

#include "_ref/gen/frame_inline.inc"     // WorldToCameraPoint, PrepareStereoCandidates (include/Frame.h:179-190, 230-263)

    void AssignFeaturesToGrid();
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                          const int maxLevel = -1) const;
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    void ComputeStereoMatches();
    int ComputeStereoMatches_Undistorted(bool isOnline = false);

    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    float mbf, mb;
    int N;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn, mvKeysRightUn;
    std::vector<float> mvuRight, mvDepth;
    std::vector<bool> mvStereoMatched;
    std::vector<std::vector<size_t> > mvRowIndices;
    std::vector<std::pair<int, int> > mvDistIdx;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<int> mvpMatchScore;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    cv::Mat mRcw, mtcw;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
};

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10);

    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

protected:
    float RadiusByViewingCos(const float& viewCos);
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2
