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
#include <mutex>
#include <set>
#include <utility>
#include <vector>

#include "ORBextractor.h"   // the reference's own header (-I/root/reference/include)

// include/Frame.h:46, :92-93
#define DISPARITY_THRES 50.0
#define FRAME_GRID_COLS 64
#define FRAME_GRID_ROWS 48

// Thirdparty/DBoW2/DBoW2/FeatureVector.h: class FeatureVector : public std::map<NodeId, std::vector<unsigned int> >
namespace DBoW2 {
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;
}

namespace ORB_SLAM2 {

class KeyFrame;
class Frame;

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
    cv::Mat GetNormal() { return normal_.clone(); }
    int GetIndexInKeyFrame(KeyFrame* pKF) {            // src/MapPoint.cc:470-477: mObservations.count(pKF) ? mObservations[pKF] : -1
        std::map<KeyFrame*, size_t>::iterator it = obs_.find(pKF);
        return it == obs_.end() ? -1 : (int)it->second;
    }
    bool IsInKeyFrame(KeyFrame* pKF) { return obs_.count(pKF) != 0; }                 // src/MapPoint.cc:479-483
    // the map surgery at the end of ORBmatcher::Fuse (:1061-1081) is recorded, not performed: {kind, this point, other / key}
    struct FuseEvent { int kind, point, other; };       // kind 0: AddObservation(pKF, other = idx); 1: this->Replace(other point)
    static std::vector<FuseEvent>& fuse_log() { static thread_local std::vector<FuseEvent> log; return log; }
    void AddObservation(KeyFrame* pKF, size_t idx) { obs_[pKF] = idx; fuse_log().push_back(FuseEvent{0, tag_, (int)idx}); }
    void Replace(MapPoint* pMP) { fuse_log().push_back(FuseEvent{1, tag_, pMP->tag_}); }
    // lifted verbatim from src/MapPoint.cc (oracle/_ref/gen/mappoint_methods.inc)
    void ComputeDistinctiveDescriptors();
    float GetMinDistanceInvariance();
    float GetMaxDistanceInvariance();
    int PredictScale(const float& currentDist, KeyFrame* pKF);
    int PredictScale(const float& currentDist, Frame* pF);
    float mfMinDistance = 0.f, mfMaxDistance = 0.f;     // include/MapPoint.h:178-179
    bool mbBad = false;                                 // :170; what ComputeDistinctiveDescriptors reads and writes (:164-167)
    std::map<KeyFrame*, size_t> mObservations;
    cv::Mat mDescriptor;
    struct Mutex : std::mutex {        // a std::mutex that lets the stand-in live in std::vector (copies get their own)
        Mutex() {}
        Mutex(const Mutex&) {}
        Mutex& operator=(const Mutex&) { return *this; }
    } mMutexPos, mMutexFeatures;

    bool bad_ = false;
    int nobs_ = 0;
    cv::Mat desc_, world_, normal_;
    std::map<KeyFrame*, size_t> obs_;
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
    // src/Frame.cc:670-887, lifted verbatim; cv::undistortPoints is the shim's (ref_shim/cvshim.cpp)
    void UndistortKeyPointsStereo();
    void UndistortKeyPoints();
    void ComputeImageBounds(const cv::Mat& imLeft);
    void ComputeImageBoundsStereo(const cv::Mat& imLeft);

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
    // what the frame-to-frame / key-frame searches read in addition (include/Frame.h:290-293, 344, 355, 372, 382-384)
    static float fx, fy, cx, cy;
    cv::Mat mTcw;
    cv::Mat mK, mK_ori, mDistCoef, mR, mP, mK_right, mDistCoef_right, mR_right, mP_right;     // include/Frame.h:295-307, all CV_32F
    std::vector<bool> mvbOutlier;
    DBoW2::FeatureVector mFeatVec;
    int mnScaleLevels = 8;
    float mfLogScaleFactor = 0.f;
};

// include/KeyFrame.h: the members and accessors the lifted key-frame searches touch (:159-172, 211-260, 289)
class KeyFrame {
public:
    KeyFrame() : fx(0), fy(0), cx(0), cy(0), N(0), mnGridCols(FRAME_GRID_COLS), mnGridRows(FRAME_GRID_ROWS),
                 mfGridElementWidthInv(0), mfGridElementHeightInv(0), mnScaleLevels(8), mfLogScaleFactor(0),
                 mnMinX(0), mnMinY(0), mnMaxX(0), mnMaxY(0) {}
    // lifted verbatim from src/KeyFrame.cc (oracle/_ref/gen/keyframe_methods.inc)
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;
    bool IsInImage(const float& x, const float& y) const;

    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    bool isBad() { return bad_; }
    bool bad_ = false;
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }     // src/KeyFrame.cc:229-233
    cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }        // src/KeyFrame.cc:126-136
    cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }

    float fx, fy, cx, cy, mbf = 0.f;
    int N;
    std::vector<float> mvInvLevelSigma2;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    int mnGridCols, mnGridRows;
    float mfGridElementWidthInv, mfGridElementHeightInv;
    int mnScaleLevels;
    float mfLogScaleFactor;
    std::vector<float> mvScaleFactors, mvLevelSigma2;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;                 // include/KeyFrame.h:263-266: ints in the key frame
    std::vector<std::vector<std::vector<size_t> > > mGrid;
    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Tcw, Ow;
};

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono, double& numVisibleMpt);
    int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist);
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched, int th);
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo);
    int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0);
    int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                     const cv::Mat& t12, const float th);
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10);

    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

protected:
    bool CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF);
    float RadiusByViewingCos(const float& viewCos);
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2
