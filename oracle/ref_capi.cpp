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
#include "_ref/gen/frame_methods.inc"
#include "_ref/gen/matcher_methods.inc"
}  // namespace ORB_SLAM2

using ORB_SLAM2::Frame;
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
