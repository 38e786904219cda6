// Header-only C++ mirror of the reference's interface for this path, over the C ABI in gforb.h.
// Names, argument meaning and error behaviour follow the reference so its call sites compile unchanged:
//   ORB_SLAM2::ORBextractor            include/ORBextractor.h:76-168   (ctor, operator(), ComputePyramid, Get*, mvImagePyramid)
//   ORB_SLAM2::ORBmatcher::DescriptorDistance / TH_HIGH / TH_LOW        include/ORBmatcher.h:49,  src/ORBmatcher.cc:37-38
//   Frame::ComputeStereoMatches / ComputeStereoMatches_Undistorted      src/Frame.cc:889-1078, 1167-1316
//   ORBmatcher::SearchForInitialization, Frame::AssignFeaturesToGrid / GetFeaturesInArea   src/ORBmatcher.cc:520-633, src/Frame.cc:461-476,593-643
// Define GFORB_WITH_OPENCV (and have OpenCV headers) to get the cv::InputArray / cv::KeyPoint / cv::Mat
// overloads the reference's Frame.cc / Tracking.cc use; without it the same calls take plain buffers.
// Like the reference, nothing here throws for an empty image (silent no-op); CUDA/argument failures throw
// std::runtime_error carrying gforb_last_error_string() (the reference's CUDA variant exit()s instead).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../gforb.h"

#ifdef GFORB_WITH_OPENCV
#include <opencv2/core/core.hpp>
#endif

namespace gforb {

inline void check(int rc, const char* what) {
    if (rc != GFORB_OK) throw std::runtime_error(std::string(what) + ": " + gforb_last_error_string());
}

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
        : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST), minThFAST(minThFAST) {
        gforb_params p{nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST};
        check(gforb_extractor_create(&p, nullptr, &h_), "ORBextractor");
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
        umax.resize(16);
        check(gforb_extractor_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                                     mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), umax.data()), "tables");
    }
    ~ORBextractor() { gforb_extractor_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // operator()(image, mask, keypoints, descriptors) on plain buffers; mask ignored as in the reference
    void operator()(const uint8_t* image, int w, int h, size_t pitch, std::vector<gforb_keypoint>& keypoints,
                    std::vector<uint8_t>& descriptors) {
        if (!image || w <= 0 || h <= 0) return;                       // src/ORBextractor.cc:1115-1116
        const int cap = gforb_extractor_max_keypoints(h_, w, h);
        if (cap < 0) check(cap, "operator()");
        keypoints.resize(cap);
        descriptors.resize((size_t)cap * 32);
        int n = 0;
        check(gforb_extract(h_, image, w, h, pitch, keypoints.data(), descriptors.data(), cap, &n), "operator()");
        keypoints.resize(n);
        descriptors.resize((size_t)n * 32);
    }
    void ComputePyramid(const uint8_t* image, int w, int h, size_t pitch) {
        check(gforb_compute_pyramid(h_, image, w, h, pitch), "ComputePyramid");
    }

#ifdef GFORB_WITH_OPENCV
    static_assert(sizeof(cv::KeyPoint) == sizeof(gforb_keypoint), "gforb_keypoint must mirror cv::KeyPoint");
    // exact reference signature (include/ORBextractor.h:89-91)
    void operator()(cv::InputArray _image, cv::InputArray /*mask*/, std::vector<cv::KeyPoint>& _keypoints,
                    cv::OutputArray _descriptors) {
        if (_image.empty()) return;
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);
        const int cap = gforb_extractor_max_keypoints(h_, image.cols, image.rows);
        if (cap < 0) check(cap, "operator()");
        _keypoints.resize(cap);
        cv::Mat desc(cap, 32, CV_8U);
        int n = 0;
        check(gforb_extract(h_, image.data, image.cols, image.rows, image.step,
                            reinterpret_cast<gforb_keypoint*>(_keypoints.data()), desc.data, cap, &n), "operator()");
        _keypoints.resize(n);
        if (n == 0) _descriptors.release();                            // src/ORBextractor.cc:1133-1134
        else desc.rowRange(0, n).copyTo(_descriptors);
        pyramid_stale_ = true;
    }
    void ComputePyramid(cv::Mat image) {                               // include/ORBextractor.h:130-132
        ComputePyramid(image.data, image.cols, image.rows, image.step);
        pyramid_stale_ = true;
    }
    // mvImagePyramid (include/ORBextractor.h:127): host views, downloaded on demand (19-px frame included in
    // the parent buffer, ROI view returned -- same shape as the reference's Mats)
    const std::vector<cv::Mat>& ImagePyramid() {
        if (pyramid_stale_) {
            mvImagePyramid.resize(nlevels);
            for (int l = 0; l < nlevels; ++l) {
                int w, h; size_t pitch; const uint8_t* dev;
                check(gforb_pyramid_level(h_, l, &dev, &w, &h, &pitch), "mvImagePyramid");
                cv::Mat padded(h + 38, w + 38, CV_8U);
                check(gforb_download_pyramid(h_, l, padded.data, padded.step, 1), "mvImagePyramid");
                mvImagePyramid[l] = padded(cv::Rect(19, 19, w, h));
            }
            pyramid_stale_ = false;
        }
        return mvImagePyramid;
    }
    std::vector<cv::Mat> mvImagePyramid;
#endif

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    int inline GetInitThres() { return iniThFAST; }
    int inline GetMinThres() { return minThFAST; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    gforb_extractor* handle() { return h_; }

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels, iniThFAST, minThFAST;
    std::vector<int32_t> mnFeaturesPerLevel, umax;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    gforb_extractor* h_ = nullptr;
    bool pyramid_stale_ = true;
};

// The members of ORB_SLAM2::Frame that the matcher searches read (include/Frame.h:279-296,358-369): undistorted keys,
// descriptors (N x 32, row stride 32) and the image bounds the 64 x 48 feature grid spans.
struct Point2f { float x, y; };                 // layout of cv::Point2f
struct FrameView {
    const std::vector<gforb_keypoint>* mvKeysUn;
    const uint8_t* mDescriptors;
    gforb_frame_bounds bounds;                  // mnMinX, mnMaxX, mnMinY, mnMaxY
};

class ORBmatcher {
public:
    static const int TH_LOW = GFORB_TH_LOW;
    static const int TH_HIGH = GFORB_TH_HIGH;
    static const int HISTO_LENGTH = 30;
    static int DescriptorDistance(const uint8_t* a, const uint8_t* b) { return gforb_hamming256(a, b); }
#ifdef GFORB_WITH_OPENCV
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return gforb_hamming256(a.data, b.data); }
#endif
    // include/ORBmatcher.h:46
    ORBmatcher(float nnratio = 0.6f, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    ~ORBmatcher() { if (h_) gforb_matcher_destroy(h_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // src/ORBmatcher.cc:520-633.  PointT = cv::Point2f or gforb::Point2f.
    template <class PointT>
    int SearchForInitialization(const FrameView& F1, const FrameView& F2, std::vector<PointT>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10) {
        static_assert(sizeof(PointT) == 2 * sizeof(float), "vbPrevMatched must hold {float x, y}");
        const int n1 = (int)F1.mvKeysUn->size(), n2 = (int)F2.mvKeysUn->size();
        vnMatches12.assign(n1, -1);
        if ((int)vbPrevMatched.size() < n1) throw std::runtime_error("SearchForInitialization: vbPrevMatched too short");
        int nmatches = 0;
        check(gforb_search_for_initialization(handle(), &F2.bounds, F1.mvKeysUn->data(), F1.mDescriptors, n1,
                                              F2.mvKeysUn->data(), F2.mDescriptors, n2,
                                              reinterpret_cast<float*>(vbPrevMatched.data()), windowSize,
                                              mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nmatches),
              "SearchForInitialization");
        return nmatches;
    }
    // The per-frame SearchByProjection forms (src/ORBmatcher.cc:155-241, 1440-1593, 1595-1721) over queries the caller
    // has projected (one gforb_proj_query + 32 descriptor bytes per MapPoint, in the reference's iteration order; gforb.h
    // says how each form fills them).  mvuRight / occupied may be NULL.  vpMatchedQuery[k] = index of the query whose
    // MapPoint F.mvpMapPoints[k] becomes (-1: leave as is, -2: set to NULL), vMatchScore[k] = F.mvpMatchScore[k].
    //   map form:        SearchByProjection(F, ..., queries, qdesc, TH_HIGH, true,  false, ...)
    //   last-frame form: SearchByProjection(F, ..., queries, qdesc, TH_HIGH, false, mbCheckOrientation, ...)
    //   key-frame form:  SearchByProjection(F, ..., queries, qdesc, ORBdist, false, mbCheckOrientation, ...)
    int SearchByProjection(const FrameView& F, const float* mvuRight, const uint8_t* occupied,
                           const std::vector<gforb_proj_query>& queries, const uint8_t* queryDescriptors, int thDist,
                           bool ratioTest, bool checkOrientation, std::vector<int>& vpMatchedQuery,
                           std::vector<int>* vMatchScore = nullptr, std::vector<int>* vQueryKey = nullptr) {
        const int n = (int)F.mvKeysUn->size(), nq = (int)queries.size();
        vpMatchedQuery.assign(n, -1);
        std::vector<int> score(n > 0 ? n : 1), qkey(nq > 0 ? nq : 1);
        int nmatches = 0;
        check(gforb_search_by_projection(handle(), &F.bounds, F.mvKeysUn->data(), F.mDescriptors, mvuRight, occupied, n,
                                         queries.data(), queryDescriptors, nq, thDist, ratioTest ? 1 : 0,
                                         checkOrientation ? 1 : 0, vpMatchedQuery.data(), score.data(), qkey.data(), &nmatches),
              "SearchByProjection");
        if (vMatchScore) vMatchScore->assign(score.begin(), score.begin() + n);
        if (vQueryKey) vQueryKey->assign(qkey.begin(), qkey.begin() + nq);
        return nmatches;
    }
    // SearchBySim3 (src/ORBmatcher.cc:1214-1438) over queries the caller's preamble derives (Sim3 projection, IsInImage,
    // distance invariance, PredictScale: :1259-1292, :1339-1372): queries12[i1] = map point i1 of KF1 in KF2
    // (GFORB_PQ_ACTIVE or 0; levels nPredictedLevel - 1 .. nPredictedLevel; r = th * mvScaleFactors[level]), desc12 = the
    // map points' descriptors; queries21 / desc21 the other way round.  vnMatches12[i1] = agreed key of KF2 or -1 (the
    // reference then sets vpMatches12[i1] = vpMapPoints2[idx2]); returns nFound.
    int SearchBySim3(const FrameView& KF1, const FrameView& KF2, const std::vector<gforb_proj_query>& queries12,
                     const uint8_t* desc12, const std::vector<gforb_proj_query>& queries21, const uint8_t* desc21,
                     std::vector<int>& vnMatches12, int thDist = TH_HIGH) {
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)KF2.mvKeysUn->size();
        if ((int)queries12.size() != n1 || (int)queries21.size() != n2)
            throw std::invalid_argument("SearchBySim3: one query per map-point slot of each key frame");
        vnMatches12.assign(n1, -1);
        int nfound = 0;
        check(gforb_search_by_sim3(handle(), &KF1.bounds, KF1.mvKeysUn->data(), KF1.mDescriptors, n1, KF2.mvKeysUn->data(),
                                   KF2.mDescriptors, n2, queries12.data(), desc12, queries21.data(), desc21, thDist, nullptr,
                                   nullptr, vnMatches12.data(), &nfound),
              "SearchBySim3");
        return nfound;
    }
    // SearchByBoW (src/ORBmatcher.cc:270-404: valid2 = NULL, thDist = TH_LOW; :635-768: thDist = TH_LOW - 1).
    // FeatVecT = DBoW2::FeatureVector or any ordered map NodeId -> vector<unsigned>.  valid1 / valid2: one byte per
    // feature, != 0 where the MapPoint exists and is not bad (NULL = all).  vnMatches12[i] = feature of frame 2 matched
    // to feature i of frame 1 (-1 none); for the Frame form vpMapPointMatches[vnMatches12[i]] = vpMapPointsKF[i].
    template <class FeatVecT>
    int SearchByBoW(const FrameView& F1, const uint8_t* valid1, const FeatVecT& featVec1, const FrameView& F2,
                    const uint8_t* valid2, const FeatVecT& featVec2, std::vector<int>& vnMatches12, int thDist = TH_LOW) {
        std::vector<int32_t> id1, st1, ix1, id2, st2, ix2;
        flatten(featVec1, id1, st1, ix1);
        flatten(featVec2, id2, st2, ix2);
        const int n1 = (int)F1.mvKeysUn->size(), n2 = (int)F2.mvKeysUn->size();
        vnMatches12.assign(n1, -1);
        int nmatches = 0;
        check(gforb_search_by_bow(handle(), F1.mvKeysUn->data(), F1.mDescriptors, valid1, n1, id1.data(), st1.data(), ix1.data(),
                                  (int)id1.size(), F2.mvKeysUn->data(), F2.mDescriptors, valid2, n2, id2.data(), st2.data(),
                                  ix2.data(), (int)id2.size(), thDist, mbCheckOrientation ? 1 : 0, vnMatches12.data(), &nmatches),
              "SearchByBoW");
        return nmatches;
    }
    // SearchForTriangulation (src/ORBmatcher.cc:770-935).  valid: feature has no MapPoint (and is stereo if bOnlyStereo);
    // stereo: mvuRight >= 0 (NULL = monocular); ep: F12, epipole, pKF2's scale factors and level sigma^2.
    template <class FeatVecT>
    int SearchForTriangulation(const FrameView& KF1, const uint8_t* valid1, const uint8_t* stereo1, const FeatVecT& featVec1,
                               const FrameView& KF2, const uint8_t* valid2, const uint8_t* stereo2, const FeatVecT& featVec2,
                               const gforb_epipolar& ep, std::vector<std::pair<size_t, size_t> >& vMatchedPairs) {
        std::vector<int32_t> id1, st1, ix1, id2, st2, ix2;
        flatten(featVec1, id1, st1, ix1);
        flatten(featVec2, id2, st2, ix2);
        const int n1 = (int)KF1.mvKeysUn->size(), n2 = (int)KF2.mvKeysUn->size();
        std::vector<int> m12(n1 > 0 ? n1 : 1, -1);
        int nmatches = 0;
        check(gforb_search_for_triangulation(handle(), KF1.mvKeysUn->data(), KF1.mDescriptors, valid1, stereo1, n1, id1.data(),
                                             st1.data(), ix1.data(), (int)id1.size(), KF2.mvKeysUn->data(), KF2.mDescriptors, valid2,
                                             stereo2, n2, id2.data(), st2.data(), ix2.data(), (int)id2.size(), &ep, TH_LOW,
                                             mbCheckOrientation ? 1 : 0, m12.data(), &nmatches),
              "SearchForTriangulation");
        vMatchedPairs.clear();
        for (int i = 0; i < n1; ++i)                                            // :923-931
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
        return nmatches;
    }
    bool checkOrientation() const { return mbCheckOrientation; }
    gforb_matcher* handle() {
        if (!h_) check(gforb_matcher_create(mfNNratio, &h_), "ORBmatcher");
        return h_;
    }

protected:
    template <class FeatVecT>
    static void flatten(const FeatVecT& fv, std::vector<int32_t>& id, std::vector<int32_t>& start, std::vector<int32_t>& idx) {
        start.push_back(0);
        for (const auto& node : fv) {
            id.push_back((int32_t)node.first);
            for (const auto i : node.second) idx.push_back((int32_t)i);
            start.push_back((int32_t)idx.size());
        }
        if (idx.empty()) idx.push_back(0);               // keep data() non-null for empty vectors
    }
    float mfNNratio;
    bool mbCheckOrientation;
    gforb_matcher* h_ = nullptr;
};

// Frame::AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:461-476, 593-643): mGrid as one CSR.
class FrameGrid {
public:
    FrameGrid(ORBmatcher& m, const std::vector<gforb_keypoint>& mvKeysUn, const gforb_frame_bounds& b)
        : m_(m), keys_(mvKeysUn), b_(b), start_(GFORB_FRAME_GRID_COLS * GFORB_FRAME_GRID_ROWS + 1), idx_(mvKeysUn.size() + 1) {
        check(gforb_assign_features_to_grid(m_.handle(), &b_, keys_.data(), (int)keys_.size(), start_.data(), idx_.data()),
              "AssignFeaturesToGrid");
    }
    // mGrid[ix][iy]
    std::vector<size_t> cell(int ix, int iy) const {
        const int c = ix * GFORB_FRAME_GRID_ROWS + iy;
        return std::vector<size_t>(idx_.begin() + start_[c], idx_.begin() + start_[c + 1]);
    }
    std::vector<size_t> GetFeaturesInArea(float x, float y, float r, int minLevel = -1, int maxLevel = -1) const {
        gforb_area_query q{x, y, r, minLevel, maxLevel};
        std::vector<int32_t> st(2), out(keys_.size() + 1);
        check(gforb_features_in_area(m_.handle(), &b_, keys_.data(), (int)keys_.size(), &q, 1, st.data(), out.data(),
                                     (int)out.size()), "GetFeaturesInArea");
        return std::vector<size_t>(out.begin(), out.begin() + st[1]);
    }

private:
    ORBmatcher& m_;
    const std::vector<gforb_keypoint>& keys_;
    gforb_frame_bounds b_;
    std::vector<int32_t> start_, idx_;
};

// Frame::UndistortKeyPoints / UndistortKeyPointsStereo (src/Frame.cc:670-758): mvKeys -> mvKeysUn through
// cv::undistortPoints(K, distCoef, R, P).  cam.dist[0] == 0 copies the keys, like the reference.
inline void UndistortKeyPoints(ORBmatcher& m, const gforb_camera& cam, const std::vector<gforb_keypoint>& mvKeys,
                               std::vector<gforb_keypoint>& mvKeysUn) {
    mvKeysUn.resize(mvKeys.size());
    check(gforb_undistort_keypoints(m.handle(), &cam, mvKeys.data(), (int)mvKeys.size(), mvKeysUn.data()), "UndistortKeyPoints");
}
// Frame::ComputeImageBounds / ComputeImageBoundsStereo (src/Frame.cc:760-887): mnMinX, mnMaxX, mnMinY, mnMaxY
inline gforb_frame_bounds ComputeImageBounds(ORBmatcher& m, const gforb_camera& cam, int cols, int rows) {
    gforb_frame_bounds b{};
    check(gforb_compute_image_bounds(m.handle(), &cam, cols, rows, &b), "ComputeImageBounds");
    return b;
}

// The pre-processing in front of the extractor: [16-bit -> 8-bit] -> cv::remap(M1, M2, INTER_LINEAR) -> cvtColor(*2GRAY)
// (ros_stereo.cc:290-295, test_Stereo.cpp:197-210, src/Tracking.cc:366-382).  M1/M2: CV_32F maps of one camera.
// HASHING::MultiIndexHashing (include/Hashing.h:171-216), the descriptor side, on the device.  Entries are int handles
// (MapPoint::mnId); the caller keeps MapPoint::mvbHashed as one uint32_t per map point.
class MultiIndexHashing {
public:
    explicit MultiIndexHashing(unsigned feature_length = 256, unsigned table_num = 32) : h_(nullptr), tables_((int)table_num) {
        check(gforb_mih_create((int)feature_length, (int)table_num, &h_), "MultiIndexHashing");
    }
    ~MultiIndexHashing() { gforb_mih_destroy(h_); }
    MultiIndexHashing(const MultiIndexHashing&) = delete;
    MultiIndexHashing& operator=(const MultiIndexHashing&) = delete;
    unsigned getSubstringLen() const { return (unsigned)gforb_mih_substring_len(h_); }
    int updateDynamics(double e) { return gforb_mih_update_dynamics(h_, e); }
    // insert(feature) for a batch, in order: desc = n x 32 bytes, ids[i] = mnId, hashed[i] = mvbHashed as bits (updated)
    std::vector<uint8_t> insert(const uint8_t* desc, const std::vector<int32_t>& ids, std::vector<uint32_t>& hashed) {
        std::vector<uint8_t> ok(ids.size());
        hashed.resize(ids.size(), 0u);
        if (!ids.empty()) check(gforb_mih_insert(h_, desc, ids.data(), (int)ids.size(), hashed.data(), ok.data()), "MultiIndexHashing::insert");
        return ok;
    }
    // query(pDesc, candidates, nactive_tables, nretrieves) for m descriptors: candidates[q] = entry ids, table by table, newest first
    void query(const uint8_t* desc, int m, std::vector<std::vector<int32_t> >& candidates, unsigned nactive_tables = 0,
               unsigned nretrieves = 20) {
        const int stride = (int)((nactive_tables ? nactive_tables : (unsigned)tables_) * nretrieves) + 1;
        std::vector<int32_t> cand((size_t)m * stride), cnt(m > 0 ? m : 1);
        if (m > 0) check(gforb_mih_query(h_, desc, m, (int)nactive_tables, (int)nretrieves, cand.data(), stride, cnt.data()), "MultiIndexHashing::query");
        candidates.resize(m);
        for (int q = 0; q < m; ++q) candidates[q].assign(cand.begin() + (size_t)q * stride, cand.begin() + (size_t)q * stride + cnt[q]);
    }
    gforb_mih* handle() const { return h_; }

private:
    gforb_mih* h_;
    int tables_;
};

class Rectifier {
public:
    Rectifier(const float* M1, const float* M2, int w, int h) : w_(w), h_(h) { check(gforb_rectifier_create(M1, M2, w, h, &h_r), "Rectifier"); }
    ~Rectifier() { if (h_r) gforb_rectifier_destroy(h_r); }
    Rectifier(const Rectifier&) = delete;
    Rectifier& operator=(const Rectifier&) = delete;
    // dst: w x h (the maps' size) 8-bit gray, tightly packed
    void operator()(const void* src, int src_w, int src_h, size_t src_pitch, int pixel_format, std::vector<uint8_t>& dst) {
        const int dw = w_ > 0 ? w_ : src_w, dh = h_ > 0 ? h_ : src_h;
        dst.resize((size_t)dw * dh);
        check(gforb_preprocess(h_r, src, src_w, src_h, src_pitch, pixel_format, dst.data(), (size_t)dw), "remap");
    }
    gforb_rectifier* handle() { return h_r; }
private:
    int w_, h_;
    gforb_rectifier* h_r = nullptr;
};

// Frame::ComputeStereoMatches(): fills mvuRight / mvDepth from the Frame's members (src/Frame.cc:889-1078).
inline int ComputeStereoMatches(ORBextractor& left, ORBextractor& right, const std::vector<gforb_keypoint>& mvKeys,
                                const std::vector<uint8_t>& mDescriptors, const std::vector<gforb_keypoint>& mvKeysRight,
                                const std::vector<uint8_t>& mDescriptorsRight, float mbf, float mb,
                                std::vector<float>& mvuRight, std::vector<float>& mvDepth,
                                int mode = GFORB_STEREO_CLASSIC, float mnMinX = 0.f) {
    const int N = (int)mvKeys.size();
    mvuRight.assign(N, -1.0f);
    mvDepth.assign(N, -1.0f);
    gforb_stereo_params prm{mbf, mb, mode, GFORB_TH_HIGH, GFORB_TH_LOW, mnMinX};
    int nmatched = 0;
    check(gforb_stereo_match(left.handle(), right.handle(), &prm, mvKeys.data(), N, mDescriptors.data(),
                             mvKeysRight.data(), (int)mvKeysRight.size(), mDescriptorsRight.data(), nullptr, nullptr,
                             mvuRight.data(), mvDepth.data(), &nmatched), "ComputeStereoMatches");
    return nmatched;
}

}  // namespace gforb
