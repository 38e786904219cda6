// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// C ABI over the REFERENCE'S OWN CODE, built into oracle/_ref/libgforb_ref.so (git-ignored; `make -C oracle ref`):
//   * /root/reference/src/ORBextractor.cc is compiled whole and unmodified (its own TU, see oracle/Makefile);
//   * the member functions listed in oracle/ref_extract.py are lifted verbatim, by line range, from src/Frame.cc,
//     include/Frame.h and src/ORBmatcher.cc at build time and compiled HERE against oracle/ref_stub.hpp;
//   * OpenCV, which this image lacks, is stood in for by oracle/ref_shim/ (containers written here; image primitives
//     forwarded to the cv2-pinned oracle/cvprim.cpp).
// The entry points carry the same names and signatures as oracle/gforb_oracle.h so that oracle/pyoracle.py binds
// either library: tests/test_oracle_vs_ref.py runs both on the same inputs and demands identical bits.
// Everything in this file is marshalling: no search, distance, threshold or tie rule is decided here.
#include <atomic>
#include <chrono>
#include <cstring>
#include <mutex>
#include <thread>

#include "ref_shim/cvshim.hpp"
#include "ref_shim/mono_new.hpp"

using namespace std;   // the reference's Frame.h is written under `using namespace std`

#include "ref_stub.hpp"
#include "gforb_oracle.h"

namespace ORB_SLAM2 {
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
#include "_ref/gen/frame_methods.inc"
#include "_ref/gen/mappoint_methods.inc"
#include "_ref/gen/keyframe_methods.inc"
#include "_ref/gen/matcher_methods.inc"
}  // namespace ORB_SLAM2

using ORB_SLAM2::Frame;
using ORB_SLAM2::KeyFrame;
using ORB_SLAM2::MapPoint;
using ORB_SLAM2::ORBmatcher;

namespace {

// protected members of the reference class, reached through a derived class (no edit of the reference header)
struct RefExtractor : ORB_SLAM2::ORBextractor {
    RefExtractor(int nf, float sf, int nl, int a, int b) : ORB_SLAM2::ORBextractor(nf, sf, nl, a, b) {}
    using ORB_SLAM2::ORBextractor::ComputeKeyPointsOctTree;
    using ORB_SLAM2::ORBextractor::DistributeOctTree;
    using ORB_SLAM2::ORBextractor::mnFeaturesPerLevel;
    using ORB_SLAM2::ORBextractor::mvInvLevelSigma2;
    using ORB_SLAM2::ORBextractor::mvInvScaleFactor;
    using ORB_SLAM2::ORBextractor::mvLevelSigma2;
    using ORB_SLAM2::ORBextractor::mvScaleFactor;
    using ORB_SLAM2::ORBextractor::nlevels;
    using ORB_SLAM2::ORBextractor::pattern;
    using ORB_SLAM2::ORBextractor::umax;
};

// the Frame class keeps its image bounds and grid pitch in statics (include/Frame.h:367-393)
std::mutex g_frame_statics;

void set_bounds(const float* b) {
    Frame::mnMinX = b[0]; Frame::mnMaxX = b[1]; Frame::mnMinY = b[2]; Frame::mnMaxY = b[3];
    // src/Frame.cc:129-130
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / (Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / (Frame::mnMaxY - Frame::mnMinY);
}

vector<cv::KeyPoint> to_keys(const orc_keypoint* k, int n) {
    vector<cv::KeyPoint> v((size_t)(n > 0 ? n : 0));
    if (n > 0) memcpy((void*)v.data(), k, (size_t)n * sizeof(orc_keypoint));
    return v;
}

cv::Mat wrap_desc(const uint8_t* d, int n) { return n > 0 ? cv::Mat(n, 32, CV_8UC1, (void*)d, 32) : cv::Mat(); }

}  // namespace

struct orc_extractor {
    RefExtractor ex;
    cv::FastLog log;     // FAST survivors of the last cell loop (the reference keeps vToDistributeKeys local)
    bool have_pyramid = false;
    orc_extractor(int nf, float sf, int nl, int a, int b) : ex(nf, sf, nl, a, b) {}
};

static_assert(sizeof(orc_keypoint) == sizeof(cv::KeyPoint), "cv::KeyPoint layout");

// vToDistributeKeys of one level (src/ORBextractor.cc:819-828), rebuilt from the logged cv::FAST calls: a logged point
// sits at (cell origin - level ROI origin) + (x, y) in ROI pixels, and the reference stores it relative to the 16-px
// min border.  Cells the reference retried at minThFAST logged an empty first call.
static int level_candidates(const orc_extractor* h, int level, int* xyr, int cap) {
    const cv::Mat& roi = h->ex.mvImagePyramid[level];
    const uint8_t* lo = roi.data;
    const uint8_t* hi = roi.data + (size_t)roi.rows * roi.step;
    int n = 0;
    for (const cv::FastLogEntry& e : h->log.pts) {
        if (e.origin < lo || e.origin >= hi || e.step != roi.step) continue;
        const size_t off = (size_t)(e.origin - lo);
        const int cx = (int)(off % roi.step), cy = (int)(off / roi.step);
        if (n < cap) { xyr[3 * n] = cx + e.x - 16; xyr[3 * n + 1] = cy + e.y - 16; xyr[3 * n + 2] = e.score; }
        ++n;
    }
    return n;
}

extern "C" {

// 1: operator new inside the entry points below hands out monotonically increasing addresses (ref_shim/mono_new.cpp),
// which makes the reference's address-ordered quadtree ties reproducible; 0 (default): plain malloc, as shipped.
void ref_monotonic_new(int on) { refshim::monotonic_new_enable(on); }

int orc_hamming256(const uint8_t* a, const uint8_t* b) {
    return ORBmatcher::DescriptorDistance(cv::Mat(1, 32, CV_8UC1, (void*)a, 32), cv::Mat(1, 32, CV_8UC1, (void*)b, 32));
}

orc_extractor* orc_extractor_create(int nf, float sf, int nl, int ini, int mn) {
    if (nf < 0 || nl < 1 || nl > 32 || !(sf > 1.0f)) return nullptr;
    return new orc_extractor(nf, sf, nl, ini, mn);
}
void orc_extractor_destroy(orc_extractor* ex) { delete ex; }

void orc_extractor_tables(const orc_extractor* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* quota, int* umax) {
    const RefExtractor& e = h->ex;
    const size_t n = (size_t)e.nlevels;
    if (scale) memcpy(scale, e.mvScaleFactor.data(), n * 4);
    if (inv_scale) memcpy(inv_scale, e.mvInvScaleFactor.data(), n * 4);
    if (sigma2) memcpy(sigma2, e.mvLevelSigma2.data(), n * 4);
    if (inv_sigma2) memcpy(inv_sigma2, e.mvInvLevelSigma2.data(), n * 4);
    if (quota) memcpy(quota, e.mnFeaturesPerLevel.data(), n * 4);
    if (umax) memcpy(umax, e.umax.data(), 16 * 4);
}
void orc_extractor_pattern(const orc_extractor* h, int8_t* xy) {
    for (int i = 0; i < 512; ++i) { xy[2 * i] = (int8_t)h->ex.pattern[i].x; xy[2 * i + 1] = (int8_t)h->ex.pattern[i].y; }
}

int orc_compute_pyramid(orc_extractor* h, const uint8_t* img, int w, int hh, size_t pitch) {
    if (!img || w <= 0 || hh <= 0) return -1;
    h->ex.ComputePyramid(cv::Mat(hh, w, CV_8UC1, (void*)img, pitch));
    h->have_pyramid = true;
    return 0;
}
int orc_pyramid_level(const orc_extractor* h, int level, const uint8_t** ptr, int* w, int* hh, size_t* pitch) {
    if (level < 0 || level >= h->ex.nlevels || !h->have_pyramid) return -1;
    const cv::Mat& im = h->ex.mvImagePyramid[level];
    if (ptr) *ptr = im.data;
    if (w) *w = im.cols;
    if (hh) *hh = im.rows;
    if (pitch) *pitch = im.step;
    return 0;
}
int orc_detect_level(const orc_extractor* hc, int level, int* xyr, int cap) {
    orc_extractor* h = const_cast<orc_extractor*>(hc);
    if (!h->have_pyramid || level < 0 || level >= h->ex.nlevels) return -1;
    refshim::MonoScope mono;
    h->log = cv::FastLog();
    cv::g_fast_log = &h->log;
    vector<vector<cv::KeyPoint> > all;
    h->ex.ComputeKeyPointsOctTree(all);
    cv::g_fast_log = nullptr;
    return level_candidates(h, level, xyr, cap);
}
// DistributeOctTree copies whole cv::KeyPoints from input to output, so class_id carries the input index through
int orc_distribute_octree(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* keep, int cap) {
    refshim::MonoScope mono;
    RefExtractor ex(1000, 1.2f, 8, 20, 7);
    vector<cv::KeyPoint> in((size_t)n);
    for (int i = 0; i < n; ++i)
        in[i] = cv::KeyPoint((float)xyr[3 * i], (float)xyr[3 * i + 1], 7.f, -1.f, (float)xyr[3 * i + 2], 0, i);
    const vector<cv::KeyPoint> out = ex.DistributeOctTree(in, minX, maxX, minY, maxY, N, 0);
    for (int i = 0; i < (int)out.size() && i < cap; ++i) keep[i] = out[i].class_id;
    return (int)out.size();
}
int orc_extract(orc_extractor* h, const uint8_t* img, int w, int hh, size_t pitch,
                orc_keypoint* kps, uint8_t* desc, int cap, int* level_counts) {
    refshim::MonoScope mono;
    vector<cv::KeyPoint> k;
    cv::Mat d;
    h->log = cv::FastLog();
    cv::g_fast_log = &h->log;
    // (w, h) = 0 reaches the reference's own empty-image return (src/ORBextractor.cc:1115)
    h->ex(img && w > 0 && hh > 0 ? cv::Mat(hh, w, CV_8UC1, (void*)img, pitch) : cv::Mat(), cv::Mat(), k, d);
    cv::g_fast_log = nullptr;
    if (img && w > 0 && hh > 0) h->have_pyramid = true;
    const int n = (int)k.size();
    const int m = n < cap ? n : cap;
    if (kps && m > 0) memcpy(kps, k.data(), (size_t)m * sizeof(orc_keypoint));
    if (desc)
        for (int i = 0; i < m; ++i) memcpy(desc + 32 * (size_t)i, d.ptr(i), 32);
    if (level_counts) {
        for (int l = 0; l < h->ex.nlevels; ++l) level_counts[l] = 0;
        for (const cv::KeyPoint& kp : k) level_counts[kp.octave]++;
    }
    return n;
}
int orc_last_candidates(const orc_extractor* h, int level, int* xyr, int cap) {
    if (!h->have_pyramid || level < 0 || level >= h->ex.nlevels) return -1;
    return level_candidates(h, level, xyr, cap);
}

// Frame::ComputeStereoMatches (mode 0) / Frame::ComputeStereoMatches_Undistorted(false) (mode 1).
// The thresholds are constants of the reference (TH_HIGH 100, TH_LOW 50): other values are refused (-2).
// Mode 0 indexes vRowIndices without a range check (src/Frame.cc:915,936): inputs that would leave it are refused (-3)
// rather than run into undefined behaviour; nL == 0 is answered here (the reference would read vDistIdx[0] of an
// unallocated vector, src/Frame.cc:1065).  min_d / max_d (per-key windows) do not exist in the reference -- they are
// derived from map points there; see ref_stereo_match_undistorted_mp.  dbg_* are not observable and stay untouched.
static int stereo_run(const orc_extractor* exL, const orc_extractor* exR, const orc_stereo_params* p,
                      const orc_keypoint* kL, int nL, const uint8_t* dL, const orc_keypoint* kR, int nR,
                      const uint8_t* dR, const float* mp_z, float* uRight, float* depth) {
    if (p->th_high != ORBmatcher::TH_HIGH || p->th_low != ORBmatcher::TH_LOW) return -2;
    if (!exL->have_pyramid) return -1;
    Frame F;
    F.mpORBextractorLeft = const_cast<RefExtractor*>(&exL->ex);
    F.mpORBextractorRight = const_cast<RefExtractor*>(&exR->ex);
    F.mbf = p->bf;
    F.mb = p->b;
    F.N = nL;
    F.mvScaleFactors = F.mpORBextractorLeft->GetScaleFactors();          // src/Frame.cc:75-81
    F.mvInvScaleFactors = F.mpORBextractorLeft->GetInverseScaleFactors();
    F.mDescriptors = wrap_desc(dL, nL);
    F.mDescriptorsRight = wrap_desc(dR, nR);
    if (nL == 0) return 0;
    int ret = 0;
    if (p->mode == 0) {
        if (!exR->have_pyramid) return -1;
        const int nRows = exL->ex.mvImagePyramid[0].rows;
        for (int i = 0; i < nR; ++i) {
            const float r = 2.0f * F.mvScaleFactors[kR[i].octave];
            if ((int)floor(kR[i].y - r) < 0 || (int)ceil(kR[i].y + r) >= nRows) return -3;
        }
        for (int i = 0; i < nL; ++i)
            if (!(kL[i].y >= 0) || (size_t)kL[i].y >= (size_t)nRows) return -3;
        F.mvKeys = to_keys(kL, nL);
        F.mvKeysRight = to_keys(kR, nR);
        F.ComputeStereoMatches();
    } else {
        F.mvKeysUn = to_keys(kL, nL);
        F.mvKeysRightUn = to_keys(kR, nR);
        F.mvpMapPoints.assign((size_t)nL, static_cast<MapPoint*>(NULL));
        vector<MapPoint> mps;
        if (mp_z) {
            mps.resize((size_t)nL);
            for (int i = 0; i < nL; ++i) {
                if (mp_z[i] != mp_z[i]) continue;       // NaN: no map point on this key
                mps[i].world_ = cv::Mat(3, 1, CV_32F);
                mps[i].world_.at<float>(0) = 0.f; mps[i].world_.at<float>(1) = 0.f; mps[i].world_.at<float>(2) = mp_z[i];
                F.mvpMapPoints[i] = &mps[i];
            }
            F.mRcw = cv::Mat(3, 3, CV_32F);             // identity pose: Pc = Pw (include/Frame.h:182)
            F.mtcw = cv::Mat(3, 1, CV_32F);
            for (int y = 0; y < 3; ++y) {
                F.mtcw.at<float>(y) = 0.f;
                for (int x = 0; x < 3; ++x) F.mRcw.at<float>(y, x) = x == y ? 1.f : 0.f;
            }
        }
        std::lock_guard<std::mutex> lock(g_frame_statics);
        Frame::mnMinX = p->min_x;
        ret = F.ComputeStereoMatches_Undistorted(false);
    }
    memcpy(uRight, F.mvuRight.data(), (size_t)nL * 4);
    memcpy(depth, F.mvDepth.data(), (size_t)nL * 4);
    if (p->mode == 0) {
        for (int i = 0; i < nL; ++i) ret += depth[i] > 0;        // the classic form returns void; count as the oracle does
    }
    return ret;
}

int orc_stereo_match(const orc_extractor* exL, const orc_extractor* exR, const orc_stereo_params* p,
                     const orc_keypoint* kL, int nL, const uint8_t* dL, const orc_keypoint* kR, int nR,
                     const uint8_t* dR, const float* min_d, const float* max_d, float* uRight, float* depth,
                     int*, int*, int*) {
    if (min_d || max_d) return -4;
    return stereo_run(exL, exR, p, kL, nL, dL, kR, nR, dR, nullptr, uRight, depth);
}

// _Undistorted with a matched map point on some keys: mp_z[i] = depth of key i's map point in the camera frame
// (NaN = none); the reference turns it into the per-key disparity window of src/Frame.cc:1220-1231.
int ref_stereo_match_undistorted_mp(const orc_extractor* exL, const orc_extractor* exR, const orc_stereo_params* p,
                                    const orc_keypoint* kL, int nL, const uint8_t* dL, const orc_keypoint* kR, int nR,
                                    const uint8_t* dR, const float* mp_z, float* uRight, float* depth) {
    if (p->mode != 1) return -4;
    return stereo_run(exL, exR, p, kL, nL, dL, kR, nR, dR, mp_z, uRight, depth);
}

// CPU baseline driver over the reference classes; same contract as the oracle's (oracle/gforb_oracle.h).
double orc_bench_stereo(const uint8_t* images, int nimg, int w, int h, int nframes, int nthreads, int shape,
                        int nf, float sf, int nl, int ini, int mn, float bf, float b, int do_match,
                        long long* checksum) {
    const size_t isz = (size_t)w * h;
    std::atomic<long long> sum{0};
    auto frame = [&](RefExtractor& exL, RefExtractor& exR, Frame& F, int f, bool fork) {
        cv::Mat iL(h, w, CV_8UC1, (void*)(images + isz * (size_t)((2 * f) % nimg)), (size_t)w);
        cv::Mat iR(h, w, CV_8UC1, (void*)(images + isz * (size_t)((2 * f + 1) % nimg)), (size_t)w);
        if (fork) {      // src/Frame.cc:268-271
            std::thread tl([&] { exL(iL, cv::Mat(), F.mvKeys, F.mDescriptors); });
            std::thread tr([&] { exR(iR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight); });
            tl.join();
            tr.join();
        } else {
            exL(iL, cv::Mat(), F.mvKeys, F.mDescriptors);
            exR(iR, cv::Mat(), F.mvKeysRight, F.mDescriptorsRight);
        }
        F.N = (int)F.mvKeys.size();
        long long s = (long long)F.mvKeys.size() + (long long)F.mvKeysRight.size();
        if (do_match && F.N > 0) {
            F.ComputeStereoMatches();
            for (int i = 0; i < F.N; ++i) s += F.mvuRight[i] >= 0;
        }
        sum += s;
    };
    auto setup = [&](RefExtractor& exL, RefExtractor& exR, Frame& F) {
        F.mpORBextractorLeft = &exL;
        F.mpORBextractorRight = &exR;
        F.mbf = bf;
        F.mb = b;
        F.mvScaleFactors = exL.GetScaleFactors();
        F.mvInvScaleFactors = exL.GetInverseScaleFactors();
    };
    auto t0 = std::chrono::steady_clock::now();
    if (shape == 1) {
        RefExtractor exL(nf, sf, nl, ini, mn), exR(nf, sf, nl, ini, mn);
        Frame F;
        setup(exL, exR, F);
        for (int f = 0; f < nframes; ++f) frame(exL, exR, F, f, true);
    } else {
        std::atomic<int> next{0};
        auto worker = [&] {
            RefExtractor exL(nf, sf, nl, ini, mn), exR(nf, sf, nl, ini, mn);
            Frame F;
            setup(exL, exR, F);
            for (;;) {
                const int f = next.fetch_add(1);
                if (f >= nframes) break;
                frame(exL, exR, F, f, false);
            }
        };
        vector<std::thread> pool;
        for (int t = 0; t < (nthreads > 0 ? nthreads : 1); ++t) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (checksum) *checksum = sum.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

// ---- Frame grid (src/Frame.cc:461-476, 593-658)
static void fill_frame(Frame& F, const orc_keypoint* keys, const uint8_t* desc, int n) {
    F.N = n;
    F.mvKeysUn = to_keys(keys, n);
    F.mvKeys = F.mvKeysUn;
    if (desc) F.mDescriptors = wrap_desc(desc, n);
    F.AssignFeaturesToGrid();
}

int orc_assign_features_to_grid(const orc_keypoint* keys, int n, const float* bounds, int* cell_start, int* cell_idx) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Frame F;
    fill_frame(F, keys, nullptr, n);
    int k = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
        for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = k;
            for (size_t i : F.mGrid[ix][iy]) cell_idx[k++] = (int)i;
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = k;
    return k;
}

int orc_features_in_area(const orc_keypoint* keys, int n, const float* bounds, float x, float y, float r,
                         int minLevel, int maxLevel, int* out, int cap) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Frame F;
    fill_frame(F, keys, nullptr, n);
    const vector<size_t> v = F.GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (int i = 0; i < (int)v.size() && i < cap; ++i) out[i] = (int)v[i];
    return (int)v.size();
}

int orc_search_for_initialization(const orc_keypoint* k1, const uint8_t* d1, int n1, const orc_keypoint* k2,
                                  const uint8_t* d2, int n2, const float* bounds, float* prev_xy, int windowSize,
                                  float nnratio, int checkOri, int* matches12) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Frame F1, F2;
    fill_frame(F1, k1, d1, n1);
    fill_frame(F2, k2, d2, n2);
    vector<cv::Point2f> prev((size_t)n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prev_xy[2 * i], prev_xy[2 * i + 1]);
    vector<int> m12;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int nm = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; ++i) { matches12[i] = m12[i]; prev_xy[2 * i] = prev[i].x; prev_xy[2 * i + 1] = prev[i].y; }
    return nm;
}

// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>, th) (src/ORBmatcher.cc:155-241) on caller-made map points:
// per point q: in_view, bad, level (mnTrackScaleLevel), view_cos, proj x / y / xr, observations, descriptor.
// key_obs[k] > 0 marks a key whose map point has observations (skipped, :197-199).  Outputs: key_point[k] = index of
// the point now held by key k (-1 none; a key's pre-existing point is reported as -1 as well), key_dist[k] = mvpMatchScore.
int ref_search_by_projection_mp(const orc_keypoint* keys, const uint8_t* desc, const float* u_right, const int* key_obs,
                                int n, const float* bounds, const float* scale_factors, int nlevels,
                                const uint8_t* in_view, const uint8_t* bad, const int* level, const float* view_cos,
                                const float* proj_x, const float* proj_y, const float* proj_xr, const int* observations,
                                const uint8_t* qdesc, int nq, float th, float nnratio, int* key_point, int* key_dist) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Frame F;
    fill_frame(F, keys, desc, n);
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    F.mvuRight.assign((size_t)n, -1.f);
    if (u_right) F.mvuRight.assign(u_right, u_right + n);
    F.mvpMatchScore.assign((size_t)n, 0);
    vector<MapPoint> held((size_t)n);
    F.mvpMapPoints.assign((size_t)n, static_cast<MapPoint*>(NULL));
    for (int k = 0; k < n; ++k)
        if (key_obs && key_obs[k] >= 0) { held[k].nobs_ = key_obs[k]; F.mvpMapPoints[k] = &held[k]; }
    vector<MapPoint> mps((size_t)nq);
    vector<MapPoint*> ptrs((size_t)nq);
    for (int q = 0; q < nq; ++q) {
        MapPoint& m = mps[q];
        m.mbTrackInView = in_view[q] != 0;
        m.bad_ = bad[q] != 0;
        m.mnTrackScaleLevel = level[q];
        m.mTrackViewCos = view_cos[q];
        m.mTrackProjX = proj_x[q];
        m.mTrackProjY = proj_y[q];
        m.mTrackProjXR = proj_xr[q];
        m.nobs_ = observations[q];
        m.desc_ = cv::Mat(1, 32, CV_8UC1, (void*)(qdesc + 32 * (size_t)q), 32);
        m.tag_ = q;
        ptrs[q] = &m;
    }
    ORBmatcher matcher(nnratio, true);
    const int nm = matcher.SearchByProjection(F, ptrs, th);
    for (int k = 0; k < n; ++k) {
        key_point[k] = F.mvpMapPoints[k] ? F.mvpMapPoints[k]->tag_ : -1;
        key_dist[k] = F.mvpMatchScore[k];
    }
    return nm;
}

}  // extern "C"

// =====================================================================================================================
// The pose-driven searches (round 2).  Everything the reference derives from poses and map points -- projections,
// PredictScale, windows -- is derived HERE by the reference's own text from the scene the test hands over (camera,
// poses, world points, per-point distance limits); tests/test_oracle_vs_ref_matchers.py derives the same queries in
// numpy float32 for the restatement and the two must agree on every key.
// Scene conventions: cam = {fx, fy, cx, cy, mbf, mb}; poses are 4x4 row-major float Tcw; a map point p has world xyz
// wpos[3p..], descriptor pdesc[32p..], flags pflag[p] (1 = exists, 2 = bad, 4 = has observations), and distance limits
// pdist[2p] = mfMinDistance, pdist[2p + 1] = mfMaxDistance.
// =====================================================================================================================
namespace {
struct Scene {
    vector<MapPoint> pts;
    Scene(const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, const float* pdist, const float* pnormal, int np) : pts((size_t)np) {
        for (int p = 0; p < np; ++p) {
            MapPoint& m = pts[p];
            m.world_ = cv::Mat(3, 1, CV_32F);
            for (int k = 0; k < 3; ++k) m.world_.at<float>(k) = wpos[3 * p + k];
            if (pnormal) { m.normal_ = cv::Mat(3, 1, CV_32F); for (int k = 0; k < 3; ++k) m.normal_.at<float>(k) = pnormal[3 * p + k]; }
            m.desc_ = cv::Mat(1, 32, CV_8UC1, (void*)(pdesc + 32 * (size_t)p), 32);
            m.bad_ = (pflag[p] & 2) != 0;
            m.nobs_ = (pflag[p] & 4) ? 1 : 0;
            if (pdist) { m.mfMinDistance = pdist[2 * p]; m.mfMaxDistance = pdist[2 * p + 1]; }
            m.tag_ = p;
        }
    }
    MapPoint* at(int p, const uint8_t* pflag) { return (p >= 0 && (pflag[p] & 1)) ? &pts[p] : static_cast<MapPoint*>(NULL); }
};
cv::Mat pose(const float* T) {
    cv::Mat m(4, 4, CV_32F);
    for (int i = 0; i < 16; ++i) m.at<float>(i / 4, i % 4) = T[i];
    return m;
}
void set_camera(const float* cam) { Frame::fx = cam[0]; Frame::fy = cam[1]; Frame::cx = cam[2]; Frame::cy = cam[3]; }
void frame_scale(Frame& F, const float* sf, int nl) {
    F.mvScaleFactors.assign(sf, sf + nl);
    F.mnScaleLevels = nl;
    F.mfLogScaleFactor = logf(sf[1]);                   // src/Frame.cc: mfLogScaleFactor = log(mfScaleFactor)
}
// src/KeyFrame.cc:36-67: the key frame copies the frame's keys, grid and scale tables
void fill_keyframe(KeyFrame& K, const orc_keypoint* keys, const uint8_t* desc, int n, const float* cam, const float* sf, int nl) {
    Frame F;
    fill_frame(F, keys, desc, n);
    K.N = n; K.mvKeysUn = F.mvKeysUn; K.mDescriptors = F.mDescriptors;
    K.fx = cam[0]; K.fy = cam[1]; K.cx = cam[2]; K.cy = cam[3];
    K.mfGridElementWidthInv = Frame::mfGridElementWidthInv; K.mfGridElementHeightInv = Frame::mfGridElementHeightInv;
    K.mnMinX = (int)Frame::mnMinX; K.mnMinY = (int)Frame::mnMinY; K.mnMaxX = (int)Frame::mnMaxX; K.mnMaxY = (int)Frame::mnMaxY;
    K.mGrid.resize(K.mnGridCols);
    for (int i = 0; i < K.mnGridCols; ++i) {
        K.mGrid[i].resize(K.mnGridRows);
        for (int j = 0; j < K.mnGridRows; ++j) K.mGrid[i][j] = F.mGrid[i][j];
    }
    K.mvScaleFactors.assign(sf, sf + nl);
    K.mvLevelSigma2.resize(nl);
    for (int l = 0; l < nl; ++l) K.mvLevelSigma2[l] = sf[l] * sf[l];
    K.mnScaleLevels = nl;
    K.mfLogScaleFactor = logf(sf[1]);
    K.mvuRight.assign((size_t)n, -1.f);
}
void set_pose(KeyFrame& K, const float* T) {            // src/KeyFrame.cc:84-103 (SetPose): Ow = -Rcw.t() * tcw
    K.Tcw = pose(T);
    cv::Mat Rcw = K.Tcw.rowRange(0, 3).colRange(0, 3), tcw = K.Tcw.rowRange(0, 3).col(3);
    K.Ow = -Rcw.t() * tcw;
}
void featvec(DBoW2::FeatureVector& fv, const int* node_id, const int* node_start, const int* node_idx, int nn) {
    for (int a = 0; a < nn; ++a)
        fv[(unsigned)node_id[a]] = vector<unsigned>(node_idx + node_start[a], node_idx + node_start[a + 1]);
}
}  // namespace

extern "C" {

// SearchByProjection(CurrentFrame, LastFrame, th, bMono, numVisibleMpt) (src/ORBmatcher.cc:1440-1593).
// last_point[i] = map point held by key i of the last frame (-1 none), last_outlier[i]; cur_point[k] = the point key k of the
// current frame already holds (-1 none).  Out: key_point[k] = point held by key k afterwards (-1 none).
int ref_search_by_projection_last(const orc_keypoint* ckeys, const uint8_t* cdesc, const float* cu_right, const int* cur_point, int n,
                                  const orc_keypoint* lkeys, const int* last_point, const uint8_t* last_outlier, int nlast,
                                  const float* bounds, const float* cam, const float* sf, int nl, const float* Tcw, const float* Tlw,
                                  const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, int np, float th, int bMono,
                                  int checkOri, int* key_point) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    set_camera(cam);
    Scene S(wpos, pdesc, pflag, nullptr, nullptr, np);
    Frame C, L;
    fill_frame(C, ckeys, cdesc, n);
    frame_scale(C, sf, nl);
    C.mbf = cam[4]; C.mb = cam[5];
    C.mTcw = pose(Tcw);
    C.mvuRight.assign(cu_right, cu_right + n);
    C.mvpMapPoints.resize((size_t)n);
    for (int k = 0; k < n; ++k) C.mvpMapPoints[k] = S.at(cur_point[k], pflag);
    fill_frame(L, lkeys, nullptr, nlast);
    L.mTcw = pose(Tlw);
    L.mvpMapPoints.resize((size_t)nlast);
    L.mvbOutlier.resize((size_t)nlast);
    for (int i = 0; i < nlast; ++i) { L.mvpMapPoints[i] = S.at(last_point[i], pflag); L.mvbOutlier[i] = last_outlier[i] != 0; }
    ORBmatcher matcher(0.9f, checkOri != 0);
    double vis = 0;
    const int nm = matcher.SearchByProjection(C, L, th, bMono != 0, vis);
    for (int k = 0; k < n; ++k) key_point[k] = C.mvpMapPoints[k] ? C.mvpMapPoints[k]->tag_ : -1;
    return nm;
}

// SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1595-1721).  kf_point[i] = point of key i of the
// key frame; already[p] != 0: in sAlreadyFound.
int ref_search_by_projection_kf(const orc_keypoint* ckeys, const uint8_t* cdesc, const int* cur_point, int n,
                                const orc_keypoint* kkeys, const int* kf_point, int nkf, const uint8_t* already,
                                const float* bounds, const float* cam, const float* sf, int nl, const float* Tcw,
                                const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, const float* pdist, int np,
                                float th, int ORBdist, int checkOri, int* key_point) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    set_camera(cam);
    Scene S(wpos, pdesc, pflag, pdist, nullptr, np);
    Frame C;
    fill_frame(C, ckeys, cdesc, n);
    frame_scale(C, sf, nl);
    C.mTcw = pose(Tcw);
    C.mvpMapPoints.resize((size_t)n);
    for (int k = 0; k < n; ++k) C.mvpMapPoints[k] = S.at(cur_point[k], pflag);
    KeyFrame K;
    fill_keyframe(K, kkeys, nullptr, nkf, cam, sf, nl);
    K.mvpMapPoints.resize((size_t)nkf);
    for (int i = 0; i < nkf; ++i) K.mvpMapPoints[i] = S.at(kf_point[i], pflag);
    std::set<MapPoint*> found;
    for (int p = 0; p < np; ++p) if (already[p]) found.insert(&S.pts[p]);
    ORBmatcher matcher(0.9f, checkOri != 0);
    const int nm = matcher.SearchByProjection(C, &K, found, th, ORBdist);
    for (int k = 0; k < n; ++k) key_point[k] = C.mvpMapPoints[k] ? C.mvpMapPoints[k]->tag_ : -1;
    return nm;
}

// SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:406-518, loop closing).  points[j] = candidate map points in
// order; matched_in[k] = point already matched to key k (-1 none).  Out: matched[k].
int ref_search_by_projection_loop(const orc_keypoint* kkeys, const uint8_t* kdesc, int nkf, const float* bounds, const float* cam,
                                  const float* sf, int nl, const float* Scw, const int* points, int npts, const int* matched_in,
                                  const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, const float* pdist,
                                  const float* pnormal, int np, int th, int* matched) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Scene S(wpos, pdesc, pflag, pdist, pnormal, np);
    KeyFrame K;
    fill_keyframe(K, kkeys, kdesc, nkf, cam, sf, nl);
    vector<MapPoint*> vp((size_t)npts), vm((size_t)nkf);
    for (int j = 0; j < npts; ++j) vp[j] = S.at(points[j], pflag);
    for (int k = 0; k < nkf; ++k) vm[k] = S.at(matched_in[k], pflag);
    ORBmatcher matcher(0.75f, true);
    const int nm = matcher.SearchByProjection(&K, pose(Scw), vp, vm, th);
    for (int k = 0; k < nkf; ++k) matched[k] = vm[k] ? vm[k]->tag_ : -1;
    return nm;
}

// SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (:1214-1438).  kfX_point[i] = point of key i; matches_in[i1] =
// point already matched to key i1 of KF1 (-1), whose index in KF2 is its obs2[p] (-1: not observed there).
int ref_search_by_sim3(const orc_keypoint* k1, const uint8_t* d1, const int* kf1_point, int n1, const orc_keypoint* k2,
                       const uint8_t* d2, const int* kf2_point, int n2, const float* bounds, const float* cam, const float* sf, int nl,
                       const float* T1w, const float* T2w, float s12, const float* R12, const float* t12, const int* matches_in,
                       const int* obs2, const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, const float* pdist, int np,
                       float th, int* matches12) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Scene S(wpos, pdesc, pflag, pdist, nullptr, np);
    KeyFrame K1, K2;
    fill_keyframe(K1, k1, d1, n1, cam, sf, nl);
    fill_keyframe(K2, k2, d2, n2, cam, sf, nl);
    set_pose(K1, T1w); set_pose(K2, T2w);
    K1.mvpMapPoints.resize((size_t)n1); K2.mvpMapPoints.resize((size_t)n2);
    for (int i = 0; i < n1; ++i) K1.mvpMapPoints[i] = S.at(kf1_point[i], pflag);
    for (int i = 0; i < n2; ++i) K2.mvpMapPoints[i] = S.at(kf2_point[i], pflag);
    for (int p = 0; p < np; ++p) if (obs2[p] >= 0) S.pts[p].obs_[&K2] = (size_t)obs2[p];
    vector<MapPoint*> vm((size_t)n1);
    for (int i = 0; i < n1; ++i) vm[i] = S.at(matches_in[i], pflag);
    cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
    for (int i = 0; i < 9; ++i) R.at<float>(i / 3, i % 3) = R12[i];
    for (int i = 0; i < 3; ++i) t.at<float>(i) = t12[i];
    ORBmatcher matcher(0.75f, true);
    const int nf = matcher.SearchBySim3(&K1, &K2, vm, s12, R, t, th);
    for (int i = 0; i < n1; ++i) matches12[i] = vm[i] ? vm[i]->tag_ : -1;
    return nf;
}

// Fuse(pKF, vpMapPoints, th) (:937-1087).  points[j] = candidate map points in order (-1 = NULL); kf_point[k] = point key k
// holds; in_kf[p] != 0: the point is already observed by the key frame.  The surgery at the end is logged:
// events[3e..] = {kind (0 AddObservation, 1 Replace), point, key index / other point}; returns nFused, *nevents.
int ref_fuse(const orc_keypoint* kkeys, const uint8_t* kdesc, const float* ur, const int* kf_point, int nkf, const float* bounds,
             const float* cam, const float* sf, int nl, const float* Tcw, const int* points, int npts, const uint8_t* in_kf,
             const float* wpos, const uint8_t* pdesc, const uint8_t* pflag, const float* pdist, const float* pnormal, int np,
             float th, int* events, int cap, int* nevents) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    Scene S(wpos, pdesc, pflag, pdist, pnormal, np);
    KeyFrame K;
    fill_keyframe(K, kkeys, kdesc, nkf, cam, sf, nl);
    set_pose(K, Tcw);
    K.mbf = cam[4];
    K.mvuRight.assign(ur, ur + nkf);
    K.mvInvLevelSigma2.resize(nl);
    for (int l = 0; l < nl; ++l) K.mvInvLevelSigma2[l] = 1.0f / K.mvLevelSigma2[l];      // src/ORBextractor.cc:429
    K.mvpMapPoints.resize((size_t)nkf);
    for (int k = 0; k < nkf; ++k) K.mvpMapPoints[k] = S.at(kf_point[k], pflag);
    for (int p = 0; p < np; ++p) if (in_kf[p]) S.pts[p].obs_[&K] = 0;
    vector<MapPoint*> vp((size_t)npts);
    for (int j = 0; j < npts; ++j) vp[j] = S.at(points[j], pflag);
    MapPoint::fuse_log().clear();
    ORBmatcher matcher(0.6f, true);
    const int nf = matcher.Fuse(&K, vp, th);
    const vector<MapPoint::FuseEvent>& log = MapPoint::fuse_log();
    *nevents = (int)log.size();
    for (int e = 0; e < (int)log.size() && e < cap; ++e) { events[3 * e] = log[e].kind; events[3 * e + 1] = log[e].point; events[3 * e + 2] = log[e].other; }
    return nf;
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-462) for npoints map points: point p is observed by
// set_start[p + 1] - set_start[p] key frames, observation j holding descriptor desc[32 * (set_start[p] + j)] (kf_bad[...]
// != 0: that key frame is bad and skipped).  The reference walks std::map<KeyFrame*, size_t>, i.e. by key-frame address:
// the stand-in key frames live in one vector, so that order is the caller's.  Out: best[32p..] = mDescriptor (zeros when
// the reference returns early), has[p].
int ref_distinctive_descriptors(const uint8_t* desc, const int* set_start, const uint8_t* kf_bad, int npoints, uint8_t* best, uint8_t* has) {
    for (int p = 0; p < npoints; ++p) {
        const int a = set_start[p], n = set_start[p + 1] - a;
        vector<KeyFrame> kfs((size_t)(n > 0 ? n : 1));
        MapPoint mp;
        for (int j = 0; j < n; ++j) {
            kfs[j].mDescriptors = cv::Mat(1, 32, CV_8UC1, (void*)(desc + 32 * (size_t)(a + j)), 32);
            kfs[j].bad_ = kf_bad && kf_bad[a + j];
            mp.mObservations[&kfs[j]] = 0;
        }
        mp.ComputeDistinctiveDescriptors();
        has[p] = !mp.mDescriptor.empty();
        std::memset(best + 32 * (size_t)p, 0, 32);
        if (has[p]) std::memcpy(best + 32 * (size_t)p, mp.mDescriptor.ptr(0), 32);
    }
    return 0;
}

// SearchByBoW(pKF, F, vpMapPointMatches) (:270-404): kf_point[i] = point of key i of the key frame; out f_point[k].
int ref_search_by_bow_kf_f(const orc_keypoint* kk, const uint8_t* kd, const int* kf_point, int nkf, const int* id1, const int* st1,
                           const int* ix1, int nn1, const orc_keypoint* fk, const uint8_t* fd, int nf, const int* id2,
                           const int* st2, const int* ix2, int nn2, const float* bounds, const uint8_t* pdesc,
                           const uint8_t* pflag, int np, float nnratio, int checkOri, int* f_point) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    vector<float> zero((size_t)3 * (np > 0 ? np : 1), 0.f);
    Scene S(zero.data(), pdesc, pflag, nullptr, nullptr, np);
    const float cam[6] = {1, 1, 0, 0, 0, 0}, sf[2] = {1.f, 1.2f};
    KeyFrame K;
    fill_keyframe(K, kk, kd, nkf, cam, sf, 2);
    K.mvpMapPoints.resize((size_t)nkf);
    for (int i = 0; i < nkf; ++i) K.mvpMapPoints[i] = S.at(kf_point[i], pflag);
    featvec(K.mFeatVec, id1, st1, ix1, nn1);
    Frame F;
    fill_frame(F, fk, fd, nf);
    featvec(F.mFeatVec, id2, st2, ix2, nn2);
    vector<MapPoint*> vm;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int nm = matcher.SearchByBoW(&K, F, vm);
    for (int k = 0; k < nf; ++k) f_point[k] = vm[k] ? vm[k]->tag_ : -1;
    return nm;
}

// SearchByBoW(pKF1, pKF2, vpMatches12) (:635-768): out matches12[i1] = point of KF2 matched to key i1.
int ref_search_by_bow_kf_kf(const orc_keypoint* k1, const uint8_t* d1, const int* kf1_point, int n1, const int* id1, const int* st1,
                            const int* ix1, int nn1, const orc_keypoint* k2, const uint8_t* d2, const int* kf2_point, int n2,
                            const int* id2, const int* st2, const int* ix2, int nn2, const float* bounds, const uint8_t* pdesc,
                            const uint8_t* pflag, int np, float nnratio, int checkOri, int* matches12) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    vector<float> zero((size_t)3 * (np > 0 ? np : 1), 0.f);
    Scene S(zero.data(), pdesc, pflag, nullptr, nullptr, np);
    const float cam[6] = {1, 1, 0, 0, 0, 0}, sf[2] = {1.f, 1.2f};
    KeyFrame K1, K2;
    fill_keyframe(K1, k1, d1, n1, cam, sf, 2);
    fill_keyframe(K2, k2, d2, n2, cam, sf, 2);
    K1.mvpMapPoints.resize((size_t)n1); K2.mvpMapPoints.resize((size_t)n2);
    for (int i = 0; i < n1; ++i) K1.mvpMapPoints[i] = S.at(kf1_point[i], pflag);
    for (int i = 0; i < n2; ++i) K2.mvpMapPoints[i] = S.at(kf2_point[i], pflag);
    featvec(K1.mFeatVec, id1, st1, ix1, nn1);
    featvec(K2.mFeatVec, id2, st2, ix2, nn2);
    vector<MapPoint*> vm;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const int nm = matcher.SearchByBoW(&K1, &K2, vm);
    for (int i = 0; i < n1; ++i) matches12[i] = vm[i] ? vm[i]->tag_ : -1;
    return nm;
}

// SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (:770-935).  kfX_has_point[i] != 0: the key holds a map
// point (it is skipped); out pairs[2j], pairs[2j + 1]; returns the number of pairs.
int ref_search_for_triangulation(const orc_keypoint* k1, const uint8_t* d1, const uint8_t* has1, const float* ur1, int n1,
                                 const int* id1, const int* st1, const int* ix1, int nn1, const orc_keypoint* k2, const uint8_t* d2,
                                 const uint8_t* has2, const float* ur2, int n2, const int* id2, const int* st2, const int* ix2, int nn2,
                                 const float* bounds, const float* cam, const float* sf, int nl, const float* T1w, const float* T2w,
                                 const float* F12, int bOnlyStereo, int checkOri, int* pairs, int cap) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    set_bounds(bounds);
    KeyFrame K1, K2;
    fill_keyframe(K1, k1, d1, n1, cam, sf, nl);
    fill_keyframe(K2, k2, d2, n2, cam, sf, nl);
    set_pose(K1, T1w); set_pose(K2, T2w);
    MapPoint held;
    K1.mvpMapPoints.resize((size_t)n1); K2.mvpMapPoints.resize((size_t)n2);
    for (int i = 0; i < n1; ++i) K1.mvpMapPoints[i] = has1[i] ? &held : static_cast<MapPoint*>(NULL);
    for (int i = 0; i < n2; ++i) K2.mvpMapPoints[i] = has2[i] ? &held : static_cast<MapPoint*>(NULL);
    K1.mvuRight.assign(ur1, ur1 + n1); K2.mvuRight.assign(ur2, ur2 + n2);
    featvec(K1.mFeatVec, id1, st1, ix1, nn1);
    featvec(K2.mFeatVec, id2, st2, ix2, nn2);
    cv::Mat F(3, 3, CV_32F);
    for (int i = 0; i < 9; ++i) F.at<float>(i / 3, i % 3) = F12[i];
    vector<pair<size_t, size_t> > vp;
    ORBmatcher matcher(0.6f, checkOri != 0);
    const int nm = matcher.SearchForTriangulation(&K1, &K2, F, vp, bOnlyStereo != 0);
    for (int j = 0; j < (int)vp.size() && j < cap; ++j) { pairs[2 * j] = (int)vp[j].first; pairs[2 * j + 1] = (int)vp[j].second; }
    (void)nm;
    return (int)vp.size();
}


// ---- Frame::UndistortKeyPoints / UndistortKeyPointsStereo / ComputeImageBounds / ComputeImageBoundsStereo (src/Frame.cc:670-887)
// cam: K[9], dist[8], R[9], P[12] as float (the reference keeps all four CV_32F), nd = distortion coefficients in use (4, 5 or 8);
// mono uses K and dist only (cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK)).
namespace {
struct RefCam { cv::Mat K, D, R, P; };
RefCam ref_cam(const float* K9, const float* dist, int nd, const float* R9, const float* P12) {
    RefCam c;
    c.K = cv::Mat(3, 3, CV_32F); c.D = cv::Mat(nd, 1, CV_32F);
    for (int i = 0; i < 9; ++i) c.K.at<float>(i / 3, i % 3) = K9[i];
    for (int i = 0; i < nd; ++i) c.D.at<float>(i) = dist[i];
    if (R9) { c.R = cv::Mat(3, 3, CV_32F); for (int i = 0; i < 9; ++i) c.R.at<float>(i / 3, i % 3) = R9[i]; }
    if (P12) { c.P = cv::Mat(3, 4, CV_32F); for (int i = 0; i < 12; ++i) c.P.at<float>(i / 4, i % 4) = P12[i]; }
    return c;
}
void from_keys(const vector<cv::KeyPoint>& v, orc_keypoint* out) {
    for (size_t i = 0; i < v.size(); ++i) {
        out[i].x = v[i].pt.x; out[i].y = v[i].pt.y; out[i].size = v[i].size; out[i].angle = v[i].angle;
        out[i].response = v[i].response; out[i].octave = v[i].octave; out[i].class_id = v[i].class_id;
    }
}
}  // namespace

void ref_undistort_keypoints(const float* K9, const float* dist, int nd, const orc_keypoint* keys, int n, orc_keypoint* out) {
    Frame F;
    const RefCam c = ref_cam(K9, dist, nd, nullptr, nullptr);
    F.mK = c.K; F.mDistCoef = c.D;
    F.N = n; F.mvKeys = to_keys(keys, n);
    F.UndistortKeyPoints();
    from_keys(F.mvKeysUn, out);
}
void ref_undistort_keypoints_stereo(const float* KL, const float* distL, int ndL, const float* RL, const float* PL,
                                    const float* KR, const float* distR, int ndR, const float* RR, const float* PR,
                                    const orc_keypoint* keysL, int nL, const orc_keypoint* keysR, int nR,
                                    orc_keypoint* outL, orc_keypoint* outR) {
    Frame F;
    const RefCam l = ref_cam(KL, distL, ndL, RL, PL), r = ref_cam(KR, distR, ndR, RR, PR);
    F.mK_ori = l.K; F.mDistCoef = l.D; F.mR = l.R; F.mP = l.P;
    F.mK_right = r.K; F.mDistCoef_right = r.D; F.mR_right = r.R; F.mP_right = r.P;
    F.N = nL; F.mvKeys = to_keys(keysL, nL); F.mvKeysRight = to_keys(keysR, nR);
    F.UndistortKeyPointsStereo();
    from_keys(F.mvKeysUn, outL);
    from_keys(F.mvKeysRightUn, outR);
}
// out4 = mnMinX, mnMaxX, mnMinY, mnMaxY; stereo != 0: ComputeImageBoundsStereo (K_ori, dist, R, P), else ComputeImageBounds (K, dist)
void ref_compute_image_bounds(int stereo, const float* K9, const float* dist, int nd, const float* R9, const float* P12,
                              int cols, int rows, float* out4) {
    std::lock_guard<std::mutex> lock(g_frame_statics);
    Frame F;
    const RefCam c = ref_cam(K9, dist, nd, stereo ? R9 : nullptr, stereo ? P12 : nullptr);
    F.mK = c.K; F.mK_ori = c.K; F.mDistCoef = c.D; F.mR = c.R; F.mP = c.P;
    cv::Mat im(rows, cols, CV_8UC1);
    if (stereo) F.ComputeImageBoundsStereo(im); else F.ComputeImageBounds(im);
    out4[0] = Frame::mnMinX; out4[1] = Frame::mnMaxX; out4[2] = Frame::mnMinY; out4[3] = Frame::mnMaxY;
}

}  // extern "C"
