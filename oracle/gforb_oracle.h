/* ORACLE -- TEST INFRASTRUCTURE ONLY.  C ABI over the CPU restatement so tests/ (ctypes) and bench.py's
 * cpu_baseline leg can call it.  The product library (include/gforb.h) never links or loads this. */
#ifndef GFORB_ORACLE_H
#define GFORB_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, size, angle, response; int octave, class_id; } orc_keypoint; /* == cv::KeyPoint */
typedef struct orc_extractor orc_extractor;

/* primitives (cross-checked against cv2 in tests/test_oracle_cv2.py) */
int   orc_cv_round_f(float v);
int   orc_cv_round_d(double v);
float orc_fast_atan2(float y, float x);
void  orc_border_reflect101(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch, int border);
void  orc_resize_linear_8u(const uint8_t* src, int sw, int sh, size_t spitch, uint8_t* dst, int dw, int dh, size_t dpitch);
void  orc_gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch);
/* xys: cap x 3 ints (x, y, score); returns the number found (may exceed cap; only cap are written) */
int   orc_fast9_16(const uint8_t* img, int w, int h, size_t pitch, int threshold, int nonmax, int* xys, int cap);
int   orc_hamming256(const uint8_t* a, const uint8_t* b);

/* extractor */
orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
void  orc_extractor_destroy(orc_extractor* ex);
/* each out pointer may be NULL; arrays hold nlevels entries (umax: 16) */
void  orc_extractor_tables(const orc_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                           int* quota, int* umax);
void  orc_extractor_pattern(const orc_extractor* ex, int8_t* xy1024);
int   orc_compute_pyramid(orc_extractor* ex, const uint8_t* img, int w, int h, size_t pitch);
/* level geometry; ptr = ROI origin inside the padded buffer (19-px REFLECT_101 frame around it) */
int   orc_pyramid_level(const orc_extractor* ex, int level, const uint8_t** ptr, int* w, int* h, size_t* pitch);
/* FAST candidates of one level after the cell loop (coords relative to the 16-px min border); xyr: cap x 3 */
int   orc_detect_level(const orc_extractor* ex, int level, int* xyr, int cap);
/* standalone quadtree: n candidates (x,y,response triples) -> kept indices in result order */
int   orc_distribute_octree(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int* keep, int cap);
/* full operator(): returns n (<= cap written); level_counts (nlevels) optional */
int   orc_extract(orc_extractor* ex, const uint8_t* img, int w, int h, size_t pitch,
                  orc_keypoint* kps, uint8_t* desc, int cap, int* level_counts);
/* candidates kept from the last orc_extract call */
int   orc_last_candidates(const orc_extractor* ex, int level, int* xyr, int cap);

typedef struct {
    float bf, b;
    int mode;      /* 0 classic (SAD + parabola), 1 undistorted (Hamming only) */
    int th_high, th_low;
    float min_x;
} orc_stereo_params;

/* dbg_* optional (nL ints each).  min_d/max_d optional per-key windows (mode 1). */
int   orc_stereo_match(const orc_extractor* exL, const orc_extractor* exR, const orc_stereo_params* prm,
                       const orc_keypoint* kL, int nL, const uint8_t* dL,
                       const orc_keypoint* kR, int nR, const uint8_t* dR,
                       const float* min_d, const float* max_d,
                       float* uRight, float* depth, int* dbg_best_dist, int* dbg_best_idx, int* dbg_sad);

/* CPU baseline driver: runs `nframes` stereo frames (extract L + extract R [+ classic stereo match]) over
 * the given images (frame f uses images[(2f) % nimg], images[(2f+1) % nimg], all w x h, tightly packed)
 * on `nthreads` worker threads.  shape 0 = "all cores" (each worker runs whole frames),
 * shape 1 = "reference-shaped" (one frame at a time, 2 threads = one per eye, serial match; src/Frame.cc:268-282).
 * Returns wall seconds; *checksum accumulates keypoint/match counts so the work cannot be elided. */
double orc_bench_stereo(const uint8_t* images, int nimg, int w, int h, int nframes, int nthreads, int shape,
                        int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                        float bf, float b, int do_match, long long* checksum);

/* Brute-force nearest / second-nearest search exactly as the reference's matcher loops write it
 * (src/ORBmatcher.cc:562-576): bestDist = bestDist2 = INT_MAX; for j in order: dist < bestDist -> shift,
 * else dist < bestDist2 -> second; accepted when bestDist <= threshold && bestDist < (float)bestDist2 * ratio.
 * Outputs as include/gforb.h gforb_match_bruteforce.  nthreads splits the QUERIES (results independent). */
void  orc_match_bruteforce(const uint8_t* a, int nA, const uint8_t* b, int nB, float ratio, int threshold,
                           int* best_idx, int* best_dist, int* second_dist, int* match, int nthreads);

/* Frame feature grid + ORBmatcher::SearchForInitialization (oracle/frame_grid.cpp; reference src/Frame.cc:461-476,
 * 593-656, src/ORBmatcher.cc:520-633,1723-1764).  bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}; the grid is the
 * reference's 64 x 48.  cell_start: 64*48+1 ints in (ix, iy) order (ix major, as mGrid[ix][iy]); cell_idx: n ints. */
int   orc_assign_features_to_grid(const orc_keypoint* keys, int n, const float* bounds, int* cell_start, int* cell_idx);
int   orc_features_in_area(const orc_keypoint* keys, int n, const float* bounds, float x, float y, float r,
                           int minLevel, int maxLevel, int* out, int cap);
/* prev_xy: n1 x 2 floats, in/out (vbPrevMatched); matches12: n1 ints out; returns nmatches */
int   orc_search_for_initialization(const orc_keypoint* k1, const uint8_t* d1, int n1, const orc_keypoint* k2,
                                    const uint8_t* d2, int n2, const float* bounds, float* prev_xy, int windowSize,
                                    float nnratio, int checkOri, int* matches12);

/* ORBmatcher::SearchByProjection, the three per-frame forms (src/ORBmatcher.cc:155-241, 1440-1593, 1595-1721) as one
 * loop over already-projected queries; see oracle/frame_grid.cpp.  flags: 1 active, 2 blocks later queries,
 * 4 right-coordinate gate, 8 reprojection-error gate of ORBmatcher::Fuse (needs inv_level_sigma2).  u_right /
 * occupied / inv_level_sigma2 may be NULL. */
typedef struct orc_proj_query { float x, y, r, xr, angle; int32_t min_level, max_level, flags; } orc_proj_query;
int   orc_search_by_projection(const orc_keypoint* keys, const uint8_t* desc, const float* u_right,
                               const uint8_t* occupied, int n, const float* bounds, const orc_proj_query* qs,
                               const uint8_t* qdesc, int nq, int th_dist, float nnratio, int use_ratio, int checkOri,
                               const float* inv_level_sigma2, int* key_query, int* key_dist, int* query_key);

/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1214-1438) over already-projected queries: q12[i1] = map point i1 of KF1
 * in KF2 (x, y = u, v; r = radius; max_level = nPredictedLevel; flags & 1), qd12 = the map points' descriptors; q21 /
 * qd21 the other way round.  vnMatch1[n1], vnMatch2[n2], match12[n1] (agreed idx2 or -1); returns nFound. */
int   orc_search_by_sim3(const orc_keypoint* k1, const uint8_t* d1, int n1, const orc_keypoint* k2, const uint8_t* d2, int n2,
                         const float* bounds, const orc_proj_query* q12, const uint8_t* qd12, const orc_proj_query* q21,
                         const uint8_t* qd21, int th_dist, int* vnMatch1, int* vnMatch2, int* match12);

/* ORBmatcher::SearchByBoW, both forms (src/ORBmatcher.cc:270-404, 635-768) over flattened DBoW2 feature vectors
 * (node ids ascending, CSR offsets, feature indices); see oracle/frame_grid.cpp.  valid1 / valid2 may be NULL. */
int   orc_search_by_bow(const orc_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, int n1, const int* node_id1,
                        const int* node_start1, const int* node_idx1, int nn1, const orc_keypoint* k2, const uint8_t* d2,
                        const uint8_t* valid2, int n2, const int* node_id2, const int* node_start2, const int* node_idx2,
                        int nn2, int th_dist, float nnratio, int checkOri, int* matches12);

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:770-935, CheckDistEpipolarLine :251-268); see
 * oracle/frame_grid.cpp.  F12: 9 floats row-major; valid / stereo arrays may be NULL. */
int   orc_search_for_triangulation(const orc_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, const uint8_t* stereo1,
                                   int n1, const int* node_id1, const int* node_start1, const int* node_idx1, int nn1,
                                   const orc_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, const uint8_t* stereo2,
                                   int n2, const int* node_id2, const int* node_start2, const int* node_idx2, int nn2,
                                   const float* F12, float ex, float ey, const float* scale_factors,
                                   const float* level_sigma2, int th_low, int checkOri, int* matches12);

/* Pre-processing in front of the extractor (oracle/preprocess.cpp): [16-bit -> high byte] -> cv::remap(INTER_LINEAR,
 * BORDER_CONSTANT 0, CV_32F maps; NULL maps = none, then dw x dh must equal the source size) -> cvtColor(*2GRAY).
 * fmt: 0 gray8, 1 gray16, 2 RGB8, 3 BGR8, 4 RGBA8, 5 BGRA8. */
int   orc_preprocess(const void* src, int sw, int sh, size_t spitch, int fmt, const float* mapx, const float* mapy,
                     int dw, int dh, uint8_t* dst, size_t dpitch);

/* cv::undistortPoints as Frame::UndistortKeyPoints[Stereo] calls it (oracle/undistort.cpp; src/Frame.cc:670-758).
 * K9: camera matrix row-major; dist8: k1 k2 p1 p2 k3 k4 k5 k6 (unused = 0); RR = P(3x3) * R from orc_undistort_matrix
 * (R = identity and P = K for the monocular form).  xy: n x 2 floats. */
void  orc_undistort_matrix(const double* P9, const double* R9, double* RR);
void  orc_undistort_points(const float* xy_in, int n, const double* K9, const double* dist8, const double* RR, float* xy_out);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-458) over npoints descriptor sets (CSR set_start);
 * best_idx = row within the set (-1 for an empty set), best_median = its median distance */
void  orc_distinctive_descriptors(const uint8_t* desc, const int* set_start, int npoints, int* best_idx, int* best_median);

#ifdef __cplusplus
}
#endif
#endif
