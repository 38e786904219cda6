/* gforb -- B200-native ORB stereo front-end: C ABI (drop-in boundary).
 *
 * This is the only boundary of the product library (gf-orb-slam2_b200/libgforb.so): plain pointers and
 * sizes, no C++/torch/OpenCV types.  The reference (qiaozhijian/gf-orb-slam2) has no FFI layer; its boundary
 * is the C++ class ABI of ORB_SLAM2::ORBextractor plus three Frame/ORBmatcher members.  Every entry point
 * below names the reference interface it replaces (paths relative to the reference checkout).  The
 * header-only C++ mirror with the reference's own class/method names is include/gforb/ORBextractor.hpp;
 * the binding a reference maintainer adds is shown in INTEGRATION.md.
 *
 * All compute runs in hand-written sm_100a CUDA kernels.  There is no CPU fallback: without a CUDA device
 * every compute entry returns GFORB_ERR_CUDA.
 *
 * Threading (mirrors src/Frame.cc:84-87): distinct handles may be used concurrently from different
 * threads; one handle is not re-entrant.
 */
#ifndef GFORB_H
#define GFORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFORB_VERSION 1

#if defined(__GNUC__)
#define GFORB_API __attribute__((visibility("default")))
#else
#define GFORB_API
#endif

/* ---- status codes (every entry returns one; 0 = ok). The reference has no error reporting at all
 *      (empty image -> silent return, src/ORBextractor.cc:1115; asserts compiled out). ---- */
enum {
    GFORB_OK = 0,
    GFORB_ERR_ARG = -1,        /* null pointer / bad size / bad parameter */
    GFORB_ERR_CAPACITY = -2,   /* caller buffer or configured capacity too small (nothing truncated silently) */
    GFORB_ERR_CUDA = -3,       /* CUDA runtime/driver error, or no device */
    GFORB_ERR_STATE = -4,      /* call order (e.g. stereo before extraction) */
    GFORB_ERR_OVERFLOW = -5    /* a device-side working set exceeded its configured bound (see gforb_limits) */
};
GFORB_API const char* gforb_last_error_string(void);      /* thread-local, never NULL */
GFORB_API int gforb_version(void);

/* Layout-identical to cv::KeyPoint (28 bytes) with the fields ORBextractor::operator() leaves in it
 * (src/ORBextractor.cc:839-849,1164-1170): pt in level-0 pixels, size = (int)(31*scale), angle in degrees
 * [0,360), response = FAST score, octave = level, class_id = -1. */
typedef struct gforb_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} gforb_keypoint;

/* ORBextractor constructor arguments (include/ORBextractor.h:81-82; YAML keys ORBextractor.*,
 * src/Tracking.cc:240-250). */
typedef struct gforb_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;
    int32_t ini_th_fast;
    int32_t min_th_fast;
} gforb_params;

/* Device-side bounds.  Zero = library default.  Exceeding one makes the call fail with GFORB_ERR_OVERFLOW. */
typedef struct gforb_limits {
    int32_t max_candidates_per_level; /* FAST survivors one (image, level) may hand to the quadtree; default 12288 */
} gforb_limits;

/* =====================================================================================================
 * ORBextractor   (include/ORBextractor.h:76-168, src/ORBextractor.cc)
 * ===================================================================================================== */
typedef struct gforb_extractor gforb_extractor;

/* ORBextractor::ORBextractor (src/ORBextractor.cc:409-469).  Device buffers are allocated lazily on the
 * first image and re-allocated when the image size changes (the reference learns the size per call too).
 * Uses the calling thread's current CUDA device. */
GFORB_API int gforb_extractor_create(const gforb_params* params, const gforb_limits* limits_or_null, gforb_extractor** out);
/* ~ORBextractor; Tracking deletes and re-creates extractors mid-run (src/Tracking.cc:298-320). */
GFORB_API void gforb_extractor_destroy(gforb_extractor* ex);

/* GetLevels/GetScaleFactor(s)/GetInverseScaleFactors/GetScaleSigmaSquares/GetInverseScaleSigmaSquares
 * (include/ORBextractor.h:93-119) + mnFeaturesPerLevel + umax.  Any pointer may be NULL.  Arrays hold
 * nlevels entries (umax: 16).  Host-side, no device needed. */
GFORB_API int gforb_extractor_tables(const gforb_extractor* ex, float* scale, float* inv_scale, float* sigma2,
                           float* inv_sigma2, int32_t* features_per_level, int32_t* umax16);
GFORB_API int gforb_extractor_params(const gforb_extractor* ex, gforb_params* out);

/* void operator()(InputArray image, InputArray mask, vector<KeyPoint>&, OutputArray descriptors)
 * (include/ORBextractor.h:89-91, src/ORBextractor.cc:1112-1174).  image: 8-bit gray, `pitch` bytes per
 * row, HOST memory; the mask is ignored by the reference and has no parameter here.  Writes *n keypoints
 * (level-major order, within a level the quadtree list order) and n x 32 descriptor bytes.  An empty image
 * (NULL / w<=0 / h<=0) is the reference's silent no-op: *n = 0, GFORB_OK.  If cap < n nothing is truncated:
 * *n is set and GFORB_ERR_CAPACITY returned (gforb_extractor_max_keypoints gives a sufficient cap). */
GFORB_API int gforb_extract(gforb_extractor* ex, const uint8_t* image, int w, int h, size_t pitch,
                  gforb_keypoint* keypoints, uint8_t* descriptors, int cap, int* n);
GFORB_API int gforb_extractor_max_keypoints(const gforb_extractor* ex, int w, int h);

/* void ComputePyramid(cv::Mat) (include/ORBextractor.h:130-132, src/ORBextractor.cc:1176-1201), the
 * pyramid-only entry used by the precomputed-detections Frame constructor (src/Frame.cc:182-183). */
GFORB_API int gforb_compute_pyramid(gforb_extractor* ex, const uint8_t* image, int w, int h, size_t pitch);

/* public member mvImagePyramid (include/ORBextractor.h:127): level geometry and DEVICE pointer to pixel
 * (0,0) of the level; a 19-px BORDER_REFLECT_101 frame surrounds it in the same buffer.  Valid until the
 * next extract/compute_pyramid on this handle. */
GFORB_API int gforb_pyramid_level(const gforb_extractor* ex, int level, const uint8_t** dev_ptr, int* w, int* h, size_t* pitch);
/* copy one level to HOST (with_border: (w+38) x (h+38) including the frame, else w x h) */
GFORB_API int gforb_download_pyramid(gforb_extractor* ex, int level, uint8_t* host, size_t host_pitch, int with_border);

/* stage-level introspection for parity tests: FAST survivors of `level` from the last extract, in the
 * reference's emission order (src/ORBextractor.cc:791-829), as (x, y, response) int triples relative to the
 * 16-px min border; and the per-level kept-keypoint counts. */
GFORB_API int gforb_debug_candidates(gforb_extractor* ex, int level, int32_t* xyr, int cap, int* n);
/* the GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) copy of `level` the descriptors of the last extract were sampled
 * from (src/ORBextractor.cc:1153-1155), w x h bytes to HOST */
GFORB_API int gforb_debug_blurred_level(gforb_extractor* ex, int level, uint8_t* host, size_t host_pitch);
GFORB_API int gforb_debug_level_counts(gforb_extractor* ex, int32_t* counts);

/* =====================================================================================================
 * ORBmatcher::DescriptorDistance   (include/ORBmatcher.h:49, src/ORBmatcher.cc:1768-1784)
 * ===================================================================================================== */
GFORB_API int gforb_hamming256(const void* a32, const void* b32);   /* host inline-equivalent; 0..256 */
#define GFORB_TH_HIGH 100   /* ORBmatcher::TH_HIGH, src/ORBmatcher.cc:37 */
#define GFORB_TH_LOW 50     /* ORBmatcher::TH_LOW,  src/ORBmatcher.cc:38 */

/* Brute-force nearest / second-nearest search over 256-bit descriptors with the reference's acceptance tests
 * (BASELINE.json configs[3]).  This is the loop every ORBmatcher search runs per query -- best / second-best
 * update src/ORBmatcher.cc:562-571 (also :183-197, :316-332), then `bestDist <= TH` and
 * `bestDist < (float)bestDist2 * mfNNratio` (:574-576) -- taken over ALL nB train descriptors.
 * gforb_matcher mirrors ORBmatcher(float nnratio = 0.6, bool checkOri) (include/ORBmatcher.h:46): it carries
 * mfNNratio plus the device scratch and stream.  Outputs per query i (any pointer may be NULL):
 *   best_idx[i]    index of the nearest train descriptor (lowest index among equal distances, as `<` gives; -1 if nB = 0)
 *   best_dist[i]   its distance (INT32_MAX if nB = 0, the reference's initial value)
 *   second_dist[i] second smallest distance of the multiset (INT32_MAX if nB < 2)
 *   match[i]       best_idx[i] if it passes both tests, else -1
 * nB <= 4,194,304.  Host-pointer and device-pointer (asynchronous, on `cuda_stream` or the matcher's own) forms.
 * Sets of >= 2,048 descriptors a side run as an int8 contraction on the tensor cores (bits expanded to +-1, a . b = 256 - 2 *
 * distance, tcgen05.mma kind::i8; k_match_tc.cu), smaller ones on the POPC kernels: identical outputs (GFORB_NO_TC=1 in the
 * environment forces the POPC kernels, for A/B measurements). */
typedef struct gforb_matcher gforb_matcher;
GFORB_API int gforb_matcher_create(float nnratio, gforb_matcher** out);
GFORB_API void gforb_matcher_destroy(gforb_matcher* m);
GFORB_API int gforb_match_bruteforce(gforb_matcher* m, const uint8_t* desc_a, int n_a, const uint8_t* desc_b, int n_b,
                           int threshold, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match);
GFORB_API int gforb_match_bruteforce_device(gforb_matcher* m, const uint8_t* dev_desc_a, int n_a, const uint8_t* dev_desc_b,
                                  int n_b, int threshold, int32_t* dev_best_idx, int32_t* dev_best_dist,
                                  int32_t* dev_second_dist, int32_t* dev_match, void* cuda_stream);
/* wait for the last device-form call and return its device time (CUDA events around its two kernels) */
GFORB_API int gforb_matcher_sync(gforb_matcher* m, float* last_ms_or_null);
/* The same matcher for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * It takes no handle and calls no cudaMalloc / cudaFree / stream or event creation: the per-segment scratch that
 * gforb_match_bruteforce_device keeps in a cudaMalloc'ed, growing buffer lives in a fixed 4 MiB device array inside the
 * library (held by every GPU that loads it until the process exits; csrc/k_match_static.cu).  Same kernels (the POPC
 * pair of k_match.cu), same outputs as gforb_match_bruteforce_device; all pointers are device pointers (outputs may be
 * NULL); asynchronous on `cuda_stream` (NULL = the default stream); nnratio is ORBmatcher's mfNNratio.
 * Limits: n_a <= GFORB_MATCH_STATIC_MAX_QUERIES, n_b <= 4,194,304 -- beyond either: GFORB_ERR_CAPACITY, nothing launched.
 * All calls on one device share the array: calls whose kernels may overlap on the GPU (different streams without a
 * dependency) race on it -- serialise them.  Nothing needs clearing between calls. */
#define GFORB_MATCH_STATIC_MAX_QUERIES 65536
GFORB_API int gforb_match_bruteforce_static(const uint8_t* dev_desc_a, int n_a, const uint8_t* dev_desc_b, int n_b,
                                  float nnratio, int threshold, int32_t* dev_best_idx, int32_t* dev_best_dist,
                                  int32_t* dev_second_dist, int32_t* dev_match, void* cuda_stream);
/* =====================================================================================================
 * Frame feature grid and ORBmatcher::SearchForInitialization   
 *   Frame::AssignFeaturesToGrid / PosInGrid   src/Frame.cc:461-476, 645-656; mGrid include/Frame.h:369
 *   Frame::GetFeaturesInArea                  src/Frame.cc:593-643
 *   ORBmatcher::SearchForInitialization       src/ORBmatcher.cc:520-633 (+ ComputeThreeMaxima :1723-1764)
 * The grid is the reference's FRAME_GRID_COLS x FRAME_GRID_ROWS = 64 x 48 (include/Frame.h:92-93) over the image
 * bounds mnMinX..mnMaxY (src/Frame.cc:760-887; without distortion {0, cols, 0, rows}).  mGrid[ix][iy] is returned
 * as one CSR: cell_idx = key indices ordered by (ix, iy, push_back order), cell_start[ix * 48 + iy] = first entry of
 * the cell, cell_start[64 * 48] = number of keys inside the grid.  At most 8,192 keys per frame (GFORB_ERR_CAPACITY).
 * All float arithmetic is the reference's (float, un-fused); results are bit-identical, candidate order included.
 * ===================================================================================================== */
#define GFORB_FRAME_GRID_COLS 64
#define GFORB_FRAME_GRID_ROWS 48
typedef struct gforb_frame_bounds { float min_x, max_x, min_y, max_y; } gforb_frame_bounds;   /* mnMinX, mnMaxX, mnMinY, mnMaxY */
/* one GetFeaturesInArea(x, y, r, minLevel, maxLevel) call; pass -1, -1 for the reference's default levels */
typedef struct gforb_area_query { float x, y, r; int32_t min_level, max_level; } gforb_area_query;

/* Frame::AssignFeaturesToGrid.  cell_start: 64*48+1 ints, cell_idx: n ints (host). */
GFORB_API int gforb_assign_features_to_grid(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys,
                                  int n, int32_t* cell_start, int32_t* cell_idx);
/* nq independent Frame::GetFeaturesInArea calls on one frame: query q's indices (in the reference's vIndices order) are
 * out_idx[out_start[q] .. out_start[q+1]).  out_start: nq+1 ints.  GFORB_ERR_CAPACITY (nothing truncated) if the total
 * exceeds `cap`; out_start is still complete then, so the caller can size the retry. */
GFORB_API int gforb_features_in_area(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys, int n,
                           const gforb_area_query* queries, int nq, int32_t* out_start, int32_t* out_idx, int cap);
/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) with mfNNratio = the matcher's
 * ratio and mbCheckOrientation = check_orientation.  keys1/desc1 = F1.mvKeysUn / F1.mDescriptors (n1 rows), keys2/desc2
 * = F2's; the grid of F2 is built internally.  prev_matched_xy: n1 x 2 floats, in/out (vbPrevMatched: window centres
 * in, matched F2 positions out).  matches12: n1 ints out (-1 = none).  *nmatches = the reference's return value. */
GFORB_API int gforb_search_for_initialization(gforb_matcher* m, const gforb_frame_bounds* bounds,
                                    const gforb_keypoint* keys1, const uint8_t* desc1, int n1,
                                    const gforb_keypoint* keys2, const uint8_t* desc2, int n2,
                                    float* prev_matched_xy, int window_size, int check_orientation,
                                    int32_t* matches12, int* nmatches);
/* Batched device form: n_pairs independent frame pairs, one CTA each, asynchronous on `cuda_stream` (or the matcher's
 * own).  Pair p reads keys at dev_keysX + p * strideX (elements), descriptors at dev_descX + p * strideX * 32, counts
 * dev_nX[p] (<= strideX; NULL = every pair has strideX keys) -- the layout gforb_batch keeps its keypoints in.
 * dev_prev_xy: [pair][stride1][2] in/out, dev_matches12: [pair][stride1], dev_nmatches: [pair].  Descriptor arrays
 * must be 16-byte aligned.  3 launches (grid build, window scan, in-order search); gforb_matcher_sync returns their device
 * time. */
GFORB_API int gforb_search_for_initialization_batch_device(gforb_matcher* m, const gforb_frame_bounds* bounds, int n_pairs,
                                    const gforb_keypoint* dev_keys1, const uint8_t* dev_desc1, const int32_t* dev_n1, int stride1,
                                    const gforb_keypoint* dev_keys2, const uint8_t* dev_desc2, const int32_t* dev_n2, int stride2,
                                    float* dev_prev_xy, int window_size, int check_orientation,
                                    int32_t* dev_matches12, int32_t* dev_nmatches, void* cuda_stream);
/* ORBmatcher::SearchByProjection -- the reference's three per-frame forms behind one entry:
 *   SearchByProjection(Frame&, const vector<MapPoint*>&, th)                     src/ORBmatcher.cc:155-241
 *   SearchByProjection(CurrentFrame, LastFrame, th, bMono)                       src/ORBmatcher.cc:1440-1593
 *   SearchByProjection(CurrentFrame, KeyFrame*, sAlreadyFound, th, ORBdist)      src/ORBmatcher.cc:1595-1721
 *   SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)  (loop closing)  src/ORBmatcher.cc:406-518
 *   SearchByProjection_Budget(F, vpMapPoints, th, time_constr)                   src/ORBmatcher.cc:45-153  = the first
 *       form cut off by a wall-clock budget (:141-147); the cut is where the host CPU happened to be, so it has no
 *       device equivalent: pass the query prefix you want searched (the whole search costs ~0.1 ms here)
 * All of them walk their queries IN ORDER; per query: GetFeaturesInArea(x, y, r, minLevel, maxLevel), skip keys that
 * already hold a map point, skip keys whose mvuRight is off by more than r, best (and second-best) Hamming distance,
 * threshold, [ratio test when best and second sit on the same level], assign -- and, in the two frame forms, the
 * 30-bin rotation histogram that clears the keys of all but the three largest bins.  What differs is how a query is
 * derived from a MapPoint and a pose (projection, PredictScale, RadiusByViewingCos): that stays with the caller, who
 * hands over one gforb_proj_query per candidate, in the reference's iteration order:
 *   x, y        mTrackProjX/Y (:179) | u, v (:1481-1482, :1629-1630)
 *   r           r * F.mvScaleFactors[nPredictedLevel] (:179) | th * mvScaleFactors[level] (:1495, :1648); also the
 *               bound of the right-coordinate test (:204, :1527)
 *   xr          mTrackProjXR (:203) | u - mbf * invzc (:1525)
 *   angle       the query keypoint's angle for the rotation check (:1555, :1683)
 *   min_level, max_level   as passed to GetFeaturesInArea (:179 | :1501-1506 | :1651)
 *   flags       GFORB_PQ_ACTIVE      0 = the reference `continue`s before the window (:164-168, :1469-1489, :1617-1644)
 *               GFORB_PQ_BLOCKS      pMP->Observations() > 0: later queries skip the key this one takes (:197-199,
 *                                    :1519-1521); always set for the key-frame form (:1662-1663)
 *               GFORB_PQ_RIGHT_GATE  apply the mvuRight test (map and last-frame forms)
 * The loop-closing form is: levels nPredictedLevel - 1 .. nPredictedLevel (its in-loop level test :493-496 equals the
 * level filter of GetFeaturesInArea), every query blocking, no right gate, occupied[k] = (vpMatched[k] != NULL),
 * th_dist = TH_LOW, use_ratio = 0, check_orientation = 0.
 * Frame side: keys = mvKeysUn, desc = mDescriptors, u_right = mvuRight (NULL: monocular), occupied[k] != 0 where
 * mvpMapPoints[k] already blocks (has observations; key-frame form: is non-NULL) (NULL: none).  th_dist = TH_HIGH or
 * ORBdist, 0..255 (bestDist starts at 256 with bestIdx = -1: a larger bound indexes [-1] in the reference); use_ratio = 1 for the map form (:230, mfNNratio = the matcher's ratio).
 * Out: key_query[k] = index of the query whose map point key k holds on return (-1: untouched, -2: set to NULL by
 * the rotation check), key_dist[k] = mvpMatchScore (:235), query_key[q] = key taken by query q (-1: none or removed),
 * *nmatches = the reference's return value.  <= 8,192 keys and <= 8,192 queries per frame (GFORB_ERR_CAPACITY); a local
 * map larger than that is handed over without the points the preamble rejected (they would carry flags = 0 anyway).
 * The sequential dependence (a taken key disappears from later windows) is resolved on the device: every query's
 * eight smallest candidates are computed in parallel against the initial state, then one CTA replays the queries in
 * order, up to 256 per step, re-scanning only those whose list was used up by keys taken meanwhile -- bit-identical results. */
typedef struct gforb_proj_query { float x, y, r, xr, angle; int32_t min_level, max_level, flags; } gforb_proj_query;
enum { GFORB_PQ_ACTIVE = 1, GFORB_PQ_BLOCKS = 2, GFORB_PQ_RIGHT_GATE = 4, GFORB_PQ_REPROJ_GATE = 8 };
/* The candidate search of ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:937-1087) is the
 * same entry with GFORB_PQ_ACTIVE | GFORB_PQ_REPROJ_GATE: x, y = u, v; xr = ur; r = th * mvScaleFactors[nPredictedLevel];
 * levels nPredictedLevel - 1 .. nPredictedLevel; a candidate passes when e2 * mvInvLevelSigma2[level] <= 7.8 (key with
 * mvuRight >= 0: e2 over u, v and the right coordinate) or <= 5.99 (:1022-1046).  Queries neither block nor see
 * occupancy; query_key[q] = bestIdx for bestDist <= th_dist (TH_LOW), and the MapPoint replace / add logic (:1061-1081)
 * runs on the caller's side in query order.  The Sim3 form (:1089-1212) has no error gate: plain GFORB_PQ_ACTIVE.
 * The table is set once per matcher (pKF->mvInvLevelSigma2, 1..32 levels). */
GFORB_API int gforb_matcher_set_inv_level_sigma2(gforb_matcher* m, const float* inv_level_sigma2, int n_levels);
GFORB_API int gforb_search_by_projection(gforb_matcher* m, const gforb_frame_bounds* bounds,
                                    const gforb_keypoint* keys, const uint8_t* desc, const float* u_right,
                                    const uint8_t* occupied, int n,
                                    const gforb_proj_query* queries, const uint8_t* query_desc, int nq,
                                    int th_dist, int use_ratio, int check_orientation,
                                    int32_t* key_query, int32_t* key_dist, int32_t* query_key, int* nmatches);
/* Batched device form: frame f reads keys / u_right / occupied at + f * key_stride (descriptors * 32), queries at
 * + f * query_stride, counts dev_n[f] / dev_nq[f] (NULL = every frame has key_stride / query_stride entries); outputs
 * laid out the same way.  Descriptor arrays 16-byte aligned.  Asynchronous on `cuda_stream` (or the matcher's own);
 * 3 launches (grid build, window scan, in-order resolution); gforb_matcher_sync returns their device time. */
GFORB_API int gforb_search_by_projection_batch_device(gforb_matcher* m, const gforb_frame_bounds* bounds, int n_frames,
                                    const gforb_keypoint* dev_keys, const uint8_t* dev_desc, const float* dev_u_right,
                                    const uint8_t* dev_occupied, const int32_t* dev_n, int key_stride,
                                    const gforb_proj_query* dev_queries, const uint8_t* dev_query_desc,
                                    const int32_t* dev_nq, int query_stride,
                                    int th_dist, int use_ratio, int check_orientation,
                                    int32_t* dev_key_query, int32_t* dev_key_dist, int32_t* dev_query_key,
                                    int32_t* dev_nmatches, void* cuda_stream);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)  (src/ORBmatcher.cc:1214-1438): map points of
 * KF1 are searched in KF2 through the Sim3 and the other way round, and a pair is kept when both searches agree.
 * The caller derives the queries exactly as for the projection entry (the Sim3 / pose projection, IsInImage, the
 * distance-invariance test and PredictScale need MapPoint and KeyFrame state, :1259-1292 and :1339-1372):
 *   queries12[i1], i1 in [0, n1): map point i1 of KF1 in KF2 -- x, y = u, v; r = th * pKF2->mvScaleFactors[level];
 *       min_level, max_level = level - 1, level (:1310 keeps kp.octave in that range); flags = GFORB_PQ_ACTIVE, or 0
 *       where the reference `continue`s (no map point, already matched through vpMatches12, bad, behind the camera,
 *       outside the image or the invariance range); qdesc12[i1] = pMP->GetDescriptor()
 *   queries21[i2], qdesc21: the same for KF2's map points in KF1.
 * keys1 / desc1 = pKF1->mvKeysUn / mDescriptors (n1 = N1), keys2 / desc2 likewise; bounds = the key frames' image bounds.
 * Nothing blocks and there is no ratio or orientation test: each query takes the FIRST smallest distance of its window
 * in GetFeaturesInArea order (:1303-1322) if it is <= th_dist (TH_HIGH = 100).
 * Out: match1[i1] = vnMatch1 (key of KF2 or -1), match2[i2] = vnMatch2 (both may be NULL in the host form),
 * match12[i1] = idx2 where vnMatch2[vnMatch1[i1]] == i1 (the reference then sets vpMatches12[i1] = vpMapPoints2[idx2]),
 * else -1; *nfound = the reference's return value.  <= 8,192 keys per key frame.  7 launches; the device form is
 * asynchronous on `cuda_stream` (NULL: the matcher's), descriptor arrays 16-byte aligned. */
GFORB_API int gforb_search_by_sim3(gforb_matcher* m, const gforb_frame_bounds* bounds,
                              const gforb_keypoint* keys1, const uint8_t* desc1, int n1,
                              const gforb_keypoint* keys2, const uint8_t* desc2, int n2,
                              const gforb_proj_query* queries12, const uint8_t* qdesc12,
                              const gforb_proj_query* queries21, const uint8_t* qdesc21, int th_dist,
                              int32_t* match1, int32_t* match2, int32_t* match12, int* nfound);
GFORB_API int gforb_search_by_sim3_device(gforb_matcher* m, const gforb_frame_bounds* bounds,
                              const gforb_keypoint* dev_keys1, const uint8_t* dev_desc1, int n1,
                              const gforb_keypoint* dev_keys2, const uint8_t* dev_desc2, int n2,
                              const gforb_proj_query* dev_queries12, const uint8_t* dev_qdesc12,
                              const gforb_proj_query* dev_queries21, const uint8_t* dev_qdesc21, int th_dist,
                              int32_t* dev_match1, int32_t* dev_match2, int32_t* dev_match12, int32_t* dev_nfound,
                              void* cuda_stream);
/* ORBmatcher::SearchByBoW, both forms:
 *   SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches)        src/ORBmatcher.cc:270-404
 *   SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)       src/ORBmatcher.cc:635-768
 * Frame 1 = the key frame whose features are the queries; frame 2 = the Frame / second key frame whose features are
 * the candidates.  mFeatVec (DBoW2::FeatureVector = std::map<NodeId, std::vector<unsigned>>,
 * Thirdparty/DBoW2/DBoW2/FeatureVector.h; computing it needs the vocabulary and stays with the caller) arrives
 * flattened: node_id ascending (the map's order), node_start CSR offsets (n_nodes + 1), node_idx = the vectors' feature
 * indices in their order.  valid1[i] != 0 where the query's MapPoint exists and is not bad (:300-306, :679-684; NULL =
 * all); valid2[j] the same for the candidates of the key-frame form (:700-706; NULL for the Frame form).  th_dist is
 * inclusive, 0..255: GFORB_TH_LOW for the Frame form (`<=`, :337), GFORB_TH_LOW - 1 for the key-frame form (`<`, :718);
 * mfNNratio = the matcher's ratio; check_orientation = mbCheckOrientation.
 * Out: matches12[i] = index in frame 2 matched to feature i of frame 1, -1 = none -- vpMatches12 of the key-frame
 * form; for the Frame form vpMapPointMatches[matches12[i]] = pKF's MapPoint i (a candidate is matched at most once, so
 * the two views are each other's inverse and the rotation check removes the same pairs).  *nmatches = the reference's
 * return value.  A feature belongs to one node, so the reference's skip-once-matched rule never crosses a node: one
 * warp per node, the node's queries in order inside it -- bit-identical results.  2 launches + 2 memsets. */
GFORB_API int gforb_search_by_bow(gforb_matcher* m,
                                  const gforb_keypoint* keys1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                                  const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1, int n_nodes1,
                                  const gforb_keypoint* keys2, const uint8_t* desc2, const uint8_t* valid2, int n2,
                                  const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int n_nodes2,
                                  int th_dist, int check_orientation, int32_t* matches12, int* nmatches);
/* device-pointer form (every array on the device, descriptors 16-byte aligned), asynchronous on `cuda_stream` or the
 * matcher's own; gforb_matcher_sync returns its device time */
GFORB_API int gforb_search_by_bow_device(gforb_matcher* m,
                                  const gforb_keypoint* dev_keys1, const uint8_t* dev_desc1, const uint8_t* dev_valid1, int n1,
                                  const int32_t* dev_node_id1, const int32_t* dev_node_start1, const int32_t* dev_node_idx1, int n_nodes1,
                                  const gforb_keypoint* dev_keys2, const uint8_t* dev_desc2, const uint8_t* dev_valid2, int n2,
                                  const int32_t* dev_node_id2, const int32_t* dev_node_start2, const int32_t* dev_node_idx2, int n_nodes2,
                                  int th_dist, int check_orientation, int32_t* dev_matches12, int32_t* dev_nmatches,
                                  void* cuda_stream);
/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)   src/ORBmatcher.cc:770-935, with
 * CheckDistEpipolarLine :251-268 -- LocalMapping's search for new map points between two key frames: the node walk of
 * SearchByBoW over features WITHOUT a MapPoint, a candidate must be within TH_LOW, away from the epipole unless one
 * side is a stereo point (:866-872) and on kp1's epipolar line within 3.84 sigma^2 of its level; no ratio test, the LAST
 * of equally near candidates wins (:859), nothing is ever marked matched (the reference never sets vbMatched2), so the
 * queries are independent.  Float arithmetic is the reference's, operation for operation, un-fused.
 *   valid1[i] != 0: pKF1 feature i has no MapPoint (and is a stereo point if bOnlyStereo)  :824-831; valid2 same  :844-851
 *   stereo1[i] / stereo2[j] != 0: mvuRight >= 0  (NULL: monocular key frame)
 *   ep: F12 row-major, the epipole of camera 1 in image 2 (:776-783, from the poses), pKF2's mvScaleFactors / mvLevelSigma2
 * Out: matches12[i] = feature of key frame 2 (-1 none) = vMatches12; vMatchedPairs is its list of (i, matches12[i]). */
typedef struct gforb_epipolar {
    float F12[9];
    float ex, ey;
    int32_t n_levels;               /* 1..32 */
    float scale_factors[32];
    float level_sigma2[32];
} gforb_epipolar;
GFORB_API int gforb_search_for_triangulation(gforb_matcher* m,
                                  const gforb_keypoint* keys1, const uint8_t* desc1, const uint8_t* valid1, const uint8_t* stereo1, int n1,
                                  const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1, int n_nodes1,
                                  const gforb_keypoint* keys2, const uint8_t* desc2, const uint8_t* valid2, const uint8_t* stereo2, int n2,
                                  const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int n_nodes2,
                                  const gforb_epipolar* ep, int th_low, int check_orientation, int32_t* matches12, int* nmatches);
GFORB_API int gforb_search_for_triangulation_device(gforb_matcher* m,
                                  const gforb_keypoint* dev_keys1, const uint8_t* dev_desc1, const uint8_t* dev_valid1, const uint8_t* dev_stereo1, int n1,
                                  const int32_t* dev_node_id1, const int32_t* dev_node_start1, const int32_t* dev_node_idx1, int n_nodes1,
                                  const gforb_keypoint* dev_keys2, const uint8_t* dev_desc2, const uint8_t* dev_valid2, const uint8_t* dev_stereo2, int n2,
                                  const int32_t* dev_node_id2, const int32_t* dev_node_start2, const int32_t* dev_node_idx2, int n_nodes2,
                                  const gforb_epipolar* ep, int th_low, int check_orientation, int32_t* dev_matches12,
                                  int32_t* dev_nmatches, void* cuda_stream);
/* =====================================================================================================
 * Input pre-processing in front of ORBextractor::operator()   (SURVEY.md 8f row 3), one fused kernel:
 *   16-bit -> 8-bit   cv::imread(IMREAD_GRAYSCALE) of 16-bit PNGs keeps the high byte   test/test_Stereo.cpp:100-106
 *   rectification     cv::remap(src, dst, M1, M2, cv::INTER_LINEAR), CV_32F maps, BORDER_CONSTANT 0
 *                     Examples/ROS/GF_ORB_SLAM2/src/ros_stereo.cc:290-295; test/test_Stereo.cpp:197-210
 *   colour -> gray    cv::cvtColor(CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY)   src/Tracking.cc:366-382
 * in that order (the reference rectifies in the caller and converts inside Tracking).  Bit-identical to OpenCV's 8U
 * fixed-point paths (5 fractional bits per axis in remap, Q15 gray weights).  gforb_rectifier carries one camera's
 * maps M1 (x) and M2 (y) on the device -- what cv::initUndistortRectifyMap(..., CV_32F, M1, M2) returns, dst_w x dst_h
 * floats each, tightly packed; NULL maps = no remap (then the output has the source size).
 * ===================================================================================================== */
enum { GFORB_PIX_GRAY8 = 0, GFORB_PIX_GRAY16 = 1, GFORB_PIX_RGB8 = 2, GFORB_PIX_BGR8 = 3, GFORB_PIX_RGBA8 = 4, GFORB_PIX_BGRA8 = 5 };
typedef struct gforb_rectifier gforb_rectifier;
GFORB_API int gforb_rectifier_create(const float* map_x, const float* map_y, int dst_w, int dst_h, gforb_rectifier** out);
GFORB_API void gforb_rectifier_destroy(gforb_rectifier* r);
/* one host image in, one host 8-bit gray image out (dst_w x dst_h, or src size without maps) */
GFORB_API int gforb_preprocess(gforb_rectifier* r, const void* src, int src_w, int src_h, size_t src_pitch, int pixel_format,
                     uint8_t* dst, size_t dst_pitch);
/* n_images device-resident images (image i at dev_src + i * src_image_stride bytes) -> device gray images, e.g.
 * straight into gforb_batch_device_input's buffer; asynchronous on cuda_stream (NULL: the rectifier's own).  1 launch. */
GFORB_API int gforb_preprocess_batch_device(gforb_rectifier* r, const void* dev_src, int src_w, int src_h, size_t src_pitch,
                                  size_t src_image_stride, int pixel_format, int n_images, uint8_t* dev_dst,
                                  size_t dst_pitch, size_t dst_image_stride, void* cuda_stream);
/* wait for the last device-form call; *last_ms = its device time */
GFORB_API int gforb_rectifier_sync(gforb_rectifier* r, float* last_ms_or_null);

/* =====================================================================================================
 * Keypoint undistortion and image bounds   
 *   Frame::UndistortKeyPoints / UndistortKeyPointsStereo   src/Frame.cc:670-758
 *   Frame::ComputeImageBounds / ComputeImageBoundsStereo   src/Frame.cc:760-887
 * Both run cv::undistortPoints(mat, mat, K, distCoef, R, P) (5 fixed iterations, double precision, float in/out);
 * bit-identical to OpenCV.  gforb_camera carries what the Frame members hold: K = mK (mK_ori for the stereo form),
 * dist = mDistCoef as k1 k2 p1 p2 k3 k4 k5 k6 (unused = 0), R = mR (identity for the monocular form), P = left 3x3 of
 * mP (= K for the monocular form, src/Frame.cc:745).  dist[0] == 0 is the reference's "no distortion" switch: keys
 * are copied unchanged and the bounds are the image rectangle.
 * ===================================================================================================== */
typedef struct gforb_camera { double K[9], dist[8], R[9], P[9]; } gforb_camera;
GFORB_API int gforb_undistort_keypoints(gforb_matcher* m, const gforb_camera* cam, const gforb_keypoint* keys, int n,
                              gforb_keypoint* keys_un);
/* n_images device-resident key arrays (image i at dev_keys + i * stride, dev_counts[i] keys or `stride` if NULL), the
 * layout gforb_batch keeps its keypoints in; in place allowed; asynchronous on cuda_stream (NULL: the matcher's).  1 launch. */
GFORB_API int gforb_undistort_keypoints_batch_device(gforb_matcher* m, const gforb_camera* cam, const gforb_keypoint* dev_keys,
                                           const int32_t* dev_counts, int stride, int n_images,
                                           gforb_keypoint* dev_keys_un, void* cuda_stream);
/* mnMinX, mnMaxX, mnMinY, mnMaxY from the 8 undistorted border points (floor / ceil as the reference applies them) */
GFORB_API int gforb_compute_image_bounds(gforb_matcher* m, const gforb_camera* cam, int cols, int rows, gforb_frame_bounds* out);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:397-458), the descriptor part, for
 * n_points map points at once: point p's observed descriptors are rows set_start[p] .. set_start[p+1] of `desc`
 * (what the reference gathers into vDescriptors).  best_idx[p] = the row WITHIN the set with the least median distance
 * to the others (first such row; -1 for an empty set), best_median[p] = that median (sorted[int(0.5 * (N - 1))], the
 * row's own 0 included; INT32_MAX for an empty set).  At most 1,024 descriptors per point (GFORB_ERR_CAPACITY). */
GFORB_API int gforb_distinctive_descriptors(gforb_matcher* m, const uint8_t* desc, const int32_t* set_start, int n_points,
                                  int32_t* best_idx, int32_t* best_median);

/* device memory helpers for callers without a CUDA runtime of their own (tests, bench): cudaMalloc / cudaFree /
 * synchronous cudaMemcpy; dir 0 = host->device, 1 = device->host */
GFORB_API int gforb_device_alloc(void** dev_ptr, size_t bytes);
GFORB_API void gforb_device_free(void* dev_ptr);
GFORB_API int gforb_device_copy(void* dst, const void* src, size_t bytes, int dir);

/* =====================================================================================================
 * Frame::ComputeStereoMatches / ComputeStereoMatches_Undistorted
 * (src/Frame.cc:889-1078, :1167-1316, include/Frame.h:230-263)
 * ===================================================================================================== */
enum { GFORB_STEREO_CLASSIC = 0, GFORB_STEREO_UNDISTORTED = 1 };

typedef struct gforb_stereo_params {
    float bf;          /* Frame::mbf */
    float b;           /* Frame::mb -- the reference reads it before setting it (src/Frame.cc:920 vs :322); explicit here */
    int32_t mode;      /* GFORB_STEREO_CLASSIC: Hamming + 11x11 SAD + parabola; _UNDISTORTED: Hamming only */
    int32_t th_high;   /* 0 -> GFORB_TH_HIGH */
    int32_t th_low;    /* 0 -> GFORB_TH_LOW  */
    float min_x;       /* Frame::mnMinX, _UNDISTORTED only */
} gforb_stereo_params;

/* Operates on what the Frame holds after extraction: the two key/descriptor sets (HOST) and the two
 * extractors' pyramids (device, from the last extract on exL / exR).  Writes mvuRight / mvDepth (-1 = none)
 * for the nL left keys.  min_d/max_d: optional per-key disparity windows (mode _UNDISTORTED,
 * src/Frame.cc:1220-1231), NULL = [0, bf/b).  *n_matched: keys that keep a depth (classic) or the
 * reference's return value (_Undistorted). */
GFORB_API int gforb_stereo_match(gforb_extractor* exL, gforb_extractor* exR, const gforb_stereo_params* prm,
                       const gforb_keypoint* kL, int nL, const uint8_t* dL,
                       const gforb_keypoint* kR, int nR, const uint8_t* dR,
                       const float* min_d, const float* max_d,
                       float* u_right, float* depth, int* n_matched);
/* per-left-key intermediates of the last gforb_stereo_match on exL: best Hamming (th_high if none), its
 * right index (-1), best SAD (-1 if not refined/accepted).  Any pointer may be NULL. */
GFORB_API int gforb_debug_stereo(gforb_extractor* exL, int nL, int32_t* best_dist, int32_t* best_idx, int32_t* sad);

/* =====================================================================================================
 * Batched stereo front-end: Frame::Frame(imLeft, imRight, ...) for many frames at once
 * (src/Frame.cc:248-326 = ExtractORB x2 threads + ComputeStereoMatches), never leaving the device between
 * extraction and matching.  This is the throughput path; one batch = up to max_frames stereo frames of one
 * fixed size, images [2f] = left, [2f+1] = right.
 * ===================================================================================================== */
typedef struct gforb_batch gforb_batch;

GFORB_API int gforb_batch_create(const gforb_params* params, const gforb_limits* limits_or_null,
                       int w, int h, int max_frames, gforb_batch** out);
GFORB_API void gforb_batch_destroy(gforb_batch* b);
/* run on a caller-owned cudaStream_t (e.g. torch's current stream); NULL restores the batch's own stream */
GFORB_API int gforb_batch_set_stream(gforb_batch* b, void* cuda_stream);
GFORB_API int gforb_batch_capacity(const gforb_batch* b, int* max_frames, int* max_keypoints_per_image);

/* device-resident inputs: pointer to image 0, bytes between images and between rows (tightly packed w) */
GFORB_API int gforb_batch_device_input(gforb_batch* b, uint8_t** dev_ptr, size_t* image_stride, size_t* pitch);
/* run from caller-owned device memory instead (e.g. frames another kernel wrote: a decoder, gforb_preprocess_batch_device):
 * the same layout as the batch's own buffer -- image 2f = left eye of frame f, 2f+1 = right eye, w*h tight bytes each,
 * at least 2*n_frames images for the n_frames later given to gforb_batch_run; read-only, must stay valid until the run
 * has finished.  NULL returns to the batch's own buffer.  While set, the host-input entries (gforb_batch_upload,
 * gforb_stereo_frames_batch[_async]) return GFORB_ERR_ARG. */
GFORB_API int gforb_batch_set_device_input(gforb_batch* b, const uint8_t* dev_images_or_null);
/* async H2D of n_frames stereo pairs from (ideally pinned) host memory; left/right: n_frames x h x w, tight */
GFORB_API int gforb_batch_upload(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n_frames);

/* enqueue the whole pipeline for frames [0, n_frames): pyramid, FAST, quadtree, orientation, blur,
 * descriptors for both eyes, and (prm != NULL) stereo matching.  Asynchronous. */
GFORB_API int gforb_batch_run(gforb_batch* b, int n_frames, const gforb_stereo_params* prm_or_null);

/* async D2H of results into host buffers laid out with the batch's max_keypoints_per_image (K) stride:
 * counts[2*n_frames]; keypoints[2*n_frames][K]; descriptors[2*n_frames][K][32]; u_right/depth[n_frames][K].
 * Any pointer may be NULL.  Call gforb_batch_sync before reading. */
GFORB_API int gforb_batch_download(gforb_batch* b, int n_frames, int32_t* counts, gforb_keypoint* keypoints,
                         uint8_t* descriptors, float* u_right, float* depth);
/* device-resident results, valid after gforb_batch_run on the batch's stream (or after gforb_batch_sync): the arrays
 * gforb_batch_download copies, in the same layout -- counts[2F], keypoints[2F][K], descriptors[2F][K][32] (image 2f =
 * left eye of frame f, 2f+1 = right eye), u_right / depth[F][K].  They feed the *_batch_device entries of the matcher
 * (key stride K, counts as they are) without a trip through the host.  Any out-pointer may be NULL. */
GFORB_API int gforb_batch_device_results(gforb_batch* b, const int32_t** dev_counts, const gforb_keypoint** dev_keypoints,
                               const uint8_t** dev_descriptors, const float** dev_u_right, const float** dev_depth,
                               int* max_keypoints_per_image);
/* wait for the stream and report device-side overflow flags as GFORB_ERR_OVERFLOW */
GFORB_API int gforb_batch_sync(gforb_batch* b);

/* upload + run + download + sync in one call: the end-to-end entry (host buffers in, host buffers out).
 * Calls small enough to be one pipeline chunk (< 48 frames: a live tracker's one frame at a time) are launch-bound, so
 * their whole copy -> kernels -> copy sequence is captured into a CUDA graph the first time an argument set (pointers,
 * n_frames, parameters) is seen and replayed with a single launch afterwards; up to 8 argument sets are kept per handle.
 * Pageable host buffers cannot be captured: the handle then stays on plain launches.  Results are identical. */
GFORB_API int gforb_stereo_frames_batch(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n_frames,
                              const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* keypoints,
                              uint8_t* descriptors, float* u_right, float* depth);

/* The same without the final wait: returns once everything is enqueued; outputs are complete (and the overflow flags
 * checked) when gforb_batch_sync(b) returns.  The host buffers must stay untouched until then and should be pinned
 * (gforb_host_alloc), otherwise the copies serialise.  Two batch handles used alternately (submit A, submit B, sync A,
 * submit A, ...) keep the PCIe link busy across consecutive batches: the tail of one batch (its last kernels and
 * downloads) overlaps the uploads of the next. */
GFORB_API int gforb_stereo_frames_batch_async(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n_frames,
                                    const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* keypoints,
                                    uint8_t* descriptors, float* u_right, float* depth);

/* gforb_stereo_frames_batch pipelines the batch in chunks of `frames` stereo frames (H2D of chunk c+1, kernels of
 * chunk c and D2H of chunk c-1 overlap on separate streams); 0 = automatic (16 when the batch has >= 48 frames) */
GFORB_API int gforb_batch_set_chunk(gforb_batch* b, int frames);
/* The CUDA graphs of single-chunk calls are keyed on the host POINTER VALUES, n_frames and the parameters.  A caller
 * that frees a buffer and later passes another allocation that happens to land on the same address (or changes a
 * buffer from pinned to pageable) must drop the cached graphs first; an argument set that cannot be captured (pageable
 * memory) is remembered as such and runs on plain launches without affecting the others. */
GFORB_API int gforb_batch_reset_graphs(gforb_batch* b);
/* measurement switch: gforb_stereo_frames_batch[_async] moves its chunks host -> device and the (stale) results device
 * -> host exactly as usual -- same chunks, same streams, same events -- but launches no kernel.  What that runs at is
 * the copy-only ceiling of the end-to-end path on the box at hand (bench.py reports e2e against it as e2e.link_frac). */
GFORB_API int gforb_batch_debug_copy_only(gforb_batch* b, int on);
/* kernels actually enqueued by the last gforb_batch_run / gforb_stereo_frames_batch call */
GFORB_API int gforb_batch_last_launches(const gforb_batch* b);

/* number of kernel launches one gforb_batch_run enqueues (for bench.py's gpu_launches) and per-kernel
 * algorithmic byte counts per stereo frame */
GFORB_API int gforb_batch_launches_per_run(const gforb_batch* b, int with_stereo);

/* Per-stage device timing of gforb_batch_run with CUDA events recorded on the batch's stream between the
 * stages: ms6 = {pyramid, FAST, quadtree, blur, orientation+descriptor, stereo} of the LAST run (waits for it).
 * Used by bench.py for the per-kernel roofline. */
GFORB_API int gforb_batch_stage_timing(gforb_batch* b, int enable);
GFORB_API int gforb_batch_stage_times(gforb_batch* b, float* ms6);

/* =====================================================================================================
 * Multi-index hashing of 256-bit descriptors: HASHING::MultiIndexHashing, the descriptor side
 * (include/Hashing.h:171-216, src/Hashing.cc) -- substring keys, bucket insertion and the two candidate queries, for
 * whole batches of descriptors.  Entries are caller-chosen int32 handles (MapPoint::mnId); what the reference keeps
 * inside MapPoint (mvbHashed) travels as a bit mask per entry that the caller holds.  The policy around it (which map
 * points to insert, LocalMapping.cc:619,959, Tracking.cc:1320; what to do with the candidates, Tracking.cc:3000-3060)
 * needs the map and stays with the caller.
 * ===================================================================================================== */
typedef struct gforb_mih gforb_mih;
/* MultiIndexHashing(feature_length = 256, table_num) (src/Hashing.cc:489-520).  table_num 16..32: substrings of
 * 256 / table_num = 8..16 bits, 2^bits buckets per table (the reference's own configuration is 32 tables of 8 bits,
 * src/System.cc:105).  Wider substrings (table_num < 16) are refused: GFORB_ERR_ARG. */
GFORB_API int gforb_mih_create(int feature_length, int table_num, gforb_mih** out);
GFORB_API void gforb_mih_destroy(gforb_mih* h);
GFORB_API int gforb_mih_substring_len(const gforb_mih* h);                        /* getSubstringLen (:952-955) */
/* updateDynamics(e) (:957-971): the retrieval-depth offset added to every query's count; returns the new value */
GFORB_API int gforb_mih_update_dynamics(gforb_mih* h, double e);
GFORB_API int gforb_mih_dynamics(const gforb_mih* h);
/* hash() (:53-103) for n descriptors: keys[i * table_num + t] = bucket of descriptor i in table t.  Host pointers. */
GFORB_API int gforb_mih_hash(gforb_mih* h, const uint8_t* desc, int n, uint32_t* keys);
/* insert() (:533-598) for entry[0..n) in this order: table t takes entry i unless bit t of hashed[i] is set (it is set
 * on return) or the bucket's latest entry already is entry[i]; success[i] = the reference's return value.
 * hashed == NULL: all flags clear on entry; success may be NULL.  Buckets keep their entries newest first. */
GFORB_API int gforb_mih_insert(gforb_mih* h, const uint8_t* desc, const int32_t* entry, int n, uint32_t* hashed_io, uint8_t* success);
/* device pointers (descriptors 4-byte aligned, e.g. gforb_batch_device_results' descriptor rows); dev_hashed_io and
 * dev_success are required; asynchronous on `cuda_stream` (NULL: the handle's own -- use one stream per handle);
 * 1 launch per 8,192 entries.  gforb_mih_sync waits and reports device-side overflow. */
GFORB_API int gforb_mih_insert_device(gforb_mih* h, const uint8_t* dev_desc, const int32_t* dev_entry, int n,
                                 uint32_t* dev_hashed_io, uint8_t* dev_success, void* cuda_stream);
/* query(pDesc, candidates, nactive_tables_top, nretrieves) (:699-846) for m descriptors: row q of cand (stride ints)
 * receives, table 0 .. nactive_tables - 1 in turn (0 = all tables), the newest nretrieves + dynamics entries of the
 * descriptor's bucket in that table, newest first; count[q] = how many.  stride >= nactive_tables * (nretrieves +
 * dynamics) or GFORB_ERR_CAPACITY; the unused tail of a row reads -1 (host form).  Duplicates are not removed (the
 * reference leaves that to its caller, Tracking.cc:3018-3030). */
GFORB_API int gforb_mih_query(gforb_mih* h, const uint8_t* qdesc, int m, int nactive_tables, int nretrieves,
                         int32_t* cand, int stride, int32_t* count);
GFORB_API int gforb_mih_query_device(gforb_mih* h, const uint8_t* dev_qdesc, int m, int nactive_tables, int nretrieves,
                                int32_t* dev_cand, int stride, int32_t* dev_count, void* cuda_stream);
/* query(pDesc, candidates, addi_candidates, nfeatures) (:849-950): table k is consulted when nfeatures[k] > 0 and
 * yields its newest nfeatures[k] + dynamics entries; of a run longer than NUM_CANDIDATE_FEATURES + dynamics (20 +
 * dynamics) the OLDEST 20 + dynamics go to cand and the newer ones to addi, both in table order.  nfeatures: table_num
 * ints.  stride >= sum of the per-table counts. */
GFORB_API int gforb_mih_query_split(gforb_mih* h, const uint8_t* qdesc, int m, const int32_t* nfeatures, int32_t* cand,
                               int32_t* count, int32_t* addi, int32_t* addi_count, int stride);
GFORB_API int gforb_mih_sync(gforb_mih* h);

/* pinned host memory helpers (cudaHostAlloc / cudaFreeHost) so callers can stage inputs for async copies */
GFORB_API int gforb_host_alloc(void** ptr, size_t bytes);
GFORB_API void gforb_host_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* GFORB_H */
