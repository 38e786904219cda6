// gforb internal declarations shared by the host engine and the sm_100a kernels.
#pragma once
#include <mutex>
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/gforb.h"

// NVTX stage ranges (the reference's PUSH_RANGE / POP_RANGE, include/Util_cuda.hpp:39-67, behind its USE_NVTX flag):
// built with `make NVTX=1` (-DGFORB_NVTX), every stage of Engine::run_extract and the stereo matcher is wrapped in a
// named range; the default build compiles them away.  NVTX v3 is header-only (no link dependency beyond -ldl).
#ifdef GFORB_NVTX
#include <nvtx3/nvToolsExt.h>
#define GFORB_RANGE_PUSH(name, cid)                                                                           \
    do {                                                                                                      \
        static const uint32_t gforb_nvtx_colors[] = {0x0000ff00, 0x000000ff, 0x00ffff00, 0x00ff00ff, 0x0000ffff, 0x00ff0000, 0x00ffffff}; \
        nvtxEventAttributes_t a_ = {};                                                                        \
        a_.version = NVTX_VERSION; a_.size = NVTX_EVENT_ATTRIB_STRUCT_SIZE;                                   \
        a_.colorType = NVTX_COLOR_ARGB; a_.color = gforb_nvtx_colors[(cid) % 7];                              \
        a_.messageType = NVTX_MESSAGE_TYPE_ASCII; a_.message.ascii = (name);                                  \
        nvtxRangePushEx(&a_);                                                                                 \
    } while (0)
#define GFORB_RANGE_POP() nvtxRangePop()
#else
#define GFORB_RANGE_PUSH(name, cid) do {} while (0)
#define GFORB_RANGE_POP() do {} while (0)
#endif

namespace gforb {

constexpr int kMaxLevels = 12;
constexpr int kEdge = 19;        // EDGE_THRESHOLD, src/ORBextractor.cc:74
constexpr int kResizeTileRows = 32;    // output rows of one resize CTA (4 warps x 8 rows; 64 measured slower: 0.571 vs 0.534 ms): k_pyramid.cu and the TMA box in engine.cu
constexpr int kXPad = 32;        // bytes left of the ROI in every padded row (19 of them are the frame): ROI is 32B aligned
constexpr int kMinBorder = 16;   // EDGE_THRESHOLD - 3, src/ORBextractor.cc:775
constexpr int kHalfPatch = 15;
constexpr int kMaxCellDim = 66;  // wCell < 60 always (nCols = floor(width/30)), +6 overlap

// One pyramid level of the batch arena.  All images of a level are contiguous: [level][image][row][byte].
struct LevelGeom {
    int w, h;                 // ROI size (mvImagePyramid[l].cols/rows)
    int pitch, rows;          // padded row pitch in bytes, padded row count (h + 38)
    size_t img_stride;        // bytes between consecutive images of this level
    size_t off;               // byte offset of (level, image 0, padded row 0) inside the pyramid arena
    int bpitch;               // blurred image pitch (ROI only, no frame)
    size_t bimg_stride, boff; // same for the blurred arena
    // FAST cell grid (src/ORBextractor.cc:767-808)
    int nCols, nRows, wCell, hCell, maxBX, maxBY;
    int cell_base, ncells;    // index of this level's first cell within an image's cell table
    // quadtree (src/ORBextractor.cc:539-570)
    int quota, nIni;
    float hX;
    int kcap, kbase;          // staging slots of this level within an image's staging table
    float scale, inv_scale;   // mvScaleFactor / mvInvScaleFactor
    float patch_size;         // (float)(int)(31 * scale)
    // resize tables (level >= 1): offsets into the table arena (ints)
    int tab_x, tab_y;         // start of {ofs, coef01} pairs for x (w entries) and y (h entries)
    int area2x;               // cv::resize's exact-2x INTER_AREA shortcut applies
    int tab_g;                // start of the 4-pixel-group table (8 ints per group, k_pyramid.cu), -1: generic kernel only
    int rs_bw, rs_bh;         // resize kernel: TMA box over the SOURCE level (bytes x rows) per 128x32 output tile; 0: no TMA
    int btile_base, btiles_x; // blur kernel: first warp task of this level, 128-column tasks per strip
    int bstrip_rows;          // blur kernel: output rows per warp task (balanced strips of <= 56 rows)
    int fast_bw, fast_bh;     // FAST kernel: TMA box (bytes x rows) = shared tile pitch / height for this level
};

struct Geom {
    int nlevels;
    int W, H;                 // input size
    int nimg;                 // images in the arena
    int img0;                 // first image of this launch (chunked batches); kernels index image img0 + blockIdx
    int cells_total;          // cells per image over all levels
    int cell_cap;             // survivor slots per cell (theoretical NMS maximum)
    int kstage;               // staging slots per image (sum of kcap)
    int K;                    // keypoint slots per image in the output arrays
    int ini_th, min_th;
    int max_cand;             // limit: candidates per (image, level) the quadtree kernel holds in smem
    int max_nodes;            // node table capacity of the quadtree kernel
    int btiles_total;         // blur warp tasks per image over all levels
    LevelGeom lv[kMaxLevels];
};

// TMA descriptors (kernel parameter, __grid_constant__): per-level padded pyramid as {pitch, rows, images}
struct alignas(64) TileMaps {
    CUtensorMap fast[kMaxLevels];
};

// packed FAST survivor / staged keypoint: x:12 | y:12 | response:8, coordinates relative to the 16-px border
__host__ __device__ inline uint32_t pack_cand(int x, int y, int r) {
    return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)r << 24);
}
__host__ __device__ inline int cand_x(uint32_t c) { return (int)(c & 0xFFFu); }
__host__ __device__ inline int cand_y(uint32_t c) { return (int)((c >> 12) & 0xFFFu); }
__host__ __device__ inline int cand_r(uint32_t c) { return (int)(c >> 24); }

// Device arenas of one engine
struct Buffers {
    uint8_t* in = nullptr;        // nimg x H x W
    const uint8_t* in_ext = nullptr;  // caller-owned device images in the same layout (gforb_batch_set_device_input)
    const uint8_t* input() const { return in_ext ? in_ext : in; }
    uint8_t* pyr = nullptr;       // padded pyramid arena
    uint8_t* blur = nullptr;      // blurred arena
    int32_t* tabs = nullptr;      // resize tables
    uint32_t* cell_tab = nullptr; // cells_total x 8 words: FAST per-cell geometry (k_fast.cu)
    uint32_t* cell_cnt = nullptr; // nimg x cells_total
    uint32_t* cell_pts = nullptr; // nimg x cells_total x cell_cap
    int32_t* lvl_cnt = nullptr;   // nimg x nlevels   kept keypoints per level
    uint32_t* lvl_pts = nullptr;  // nimg x kstage    staged kept keypoints (packed, level coords)
    gforb_keypoint* kps = nullptr;   // nimg x K
    uint8_t* desc = nullptr;         // nimg x K x 32
    int32_t* nkp = nullptr;          // nimg
    float* uright = nullptr;         // nframes x K
    float* depth = nullptr;          // nframes x K
    int32_t* st_best_dist = nullptr; // nframes x K
    int32_t* st_best_idx = nullptr;  // nframes x K
    int32_t* st_sad = nullptr;       // nframes x K
    int32_t* st_nmatch = nullptr;    // nframes
    int32_t* st_cut = nullptr;       // nframes x K
    int32_t* st_ticket = nullptr;    // 2 x nframes (ticket, processed count), zero between launches
    int32_t* oct_retry = nullptr;    // nimg x nlevels: quadtree pairs to redo with the full working set (k_octree.cu)
    int32_t* err = nullptr;          // [0] = sticky overflow flag bits
};

// stereo kernel view: where the two key sets and the two pyramids of frame f live
struct StereoView {
    const gforb_keypoint* kL; const gforb_keypoint* kR;   // frame 0
    const uint8_t* dL; const uint8_t* dR;
    const int32_t* nL; const int32_t* nR;
    size_t k_stride;          // keypoints between consecutive frames (same for L and R)
    size_t n_stride;          // ints between consecutive frames in nL/nR
    const uint8_t* pyrL[kMaxLevels]; const uint8_t* pyrR[kMaxLevels];  // ROI origin (0,0) of frame 0
    size_t frame_stride[kMaxLevels];  // bytes between consecutive frames in a level
    int pitch[kMaxLevels], w[kMaxLevels];
    float scale[kMaxLevels], inv_scale[kMaxLevels];
    int nRows;                // rows of level 0
    int f0;                   // first frame of this launch (chunked batches)
    const float* min_d; const float* max_d;  // optional per-key windows (frame 0), may be null
    float* uright; float* depth; int32_t* best_dist; int32_t* best_idx; int32_t* sad; int32_t* nmatch;
    int32_t* cut_dist;        // per left key: distance entering the median cut (-1 = none), out_stride layout
    int32_t* ticket; int32_t* proc_count;   // per frame, zero between launches
    size_t out_stride;        // floats between frames in the outputs
};

enum ErrBits { kErrCandOverflow = 1, kErrCellOverflow = 2, kErrNodeOverflow = 4, kErrKpOverflow = 8 };

// ---- kernel launchers (each enqueues on `s`; defined in the per-stage .cu files) ----
// images [g.img0, g.img0+nimg); rsmaps[l] = TMA map of level l-1 with level l's resize box (null: no TMA path)
// full_frames: also write the 16 outer pixels of every level's 19-px frame (one more launch); without it the levels carry the
// ROI and the 3-px frame the blur reads -- all any kernel of this library looks at
void launch_pyramid(const Geom& g, const Buffers& b, int nimg, cudaStream_t s, const CUtensorMap* rsmaps = nullptr,
                    bool full_frames = true);
int octree_small_cap(const Geom& g);
// kernels one extraction enqueues: nlevels pyramid + fast, quadtree (2 tiers when max_cand exceeds the small cap), blur, describe
inline int extract_launches(const Geom& g) { return g.nlevels + 4 + (octree_small_cap(g) < g.max_cand ? 1 : 0); }
void launch_fast(const Geom& g, const TileMaps& m, const Buffers& b, int nimg, cudaStream_t s);   // 1 launch
void launch_octree(const Geom& g, const Buffers& b, int nimg, cudaStream_t s);         // 1 launch
void launch_blur(const Geom& g, const Buffers& b, int nimg, cudaStream_t s);           // 1 launch
void launch_describe(const Geom& g, const Buffers& b, int nimg, cudaStream_t s);       // 1 launch
void launch_stereo(const StereoView& v, const gforb_stereo_params& p, int nframes, int K, cudaStream_t s);  // 1 launch
void build_fast_cell_table(const Geom& g, std::vector<uint32_t>& tab);
// brute-force matcher (k_match.cu): 2 launches
int match_segments(int nA, int nB, int sm_count);
size_t match_scratch_bytes(int nA, int nseg);
int match_max_train();
void launch_match_merge(const void* part, int nA, int nseg, float ratio, int threshold, int32_t* best_idx, int32_t* best_dist,
                        int32_t* second_dist, int32_t* match, cudaStream_t s);
// tensor-core form of the brute-force matcher (k_match_tc.cu): tcgen05.mma kind::i8 over the bits expanded to +-1
bool match_tc_eligible(int nA, int nB);
int match_tc_segments(int nA, int nB, int sm_count);
size_t match_tc_scratch_bytes(int nA, int nB, int nseg);
int launch_match_tc(const uint8_t* dA, int nA, const uint8_t* dB, int nB, float ratio, int threshold, int nseg, void* scratch,
                    int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match, cudaStream_t s);
void launch_match(const uint8_t* dA, int nA, const uint8_t* dB, int nB, float ratio, int threshold, int nseg,
                  void* scratch, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match,
                  cudaStream_t s);
// frame feature grid + SearchForInitialization (k_gridmatch.cu)
struct InitSearchArgs {
    const gforb_keypoint* k1; const uint8_t* d1; const int32_t* n1;    // per pair p: k1 + p * stride1, n1[p] (or n1_fixed)
    const gforb_keypoint* k2; const uint8_t* d2; const int32_t* n2;
    int n1_fixed, n2_fixed, stride1, stride2;
    const int32_t* cell_start; const int32_t* cell_idx;                 // frame-2 grids, one per pair
    float min_x, min_y, w_inv, h_inv;
    float* prev_xy;              // [pair][stride1][2] in/out (vbPrevMatched)
    float window, ratio;
    int check_ori, th_low;
    int32_t* matches12;          // [pair][stride1]
    int32_t* nmatches;           // [pair]
    uint4* spec;                 // [pair][stride1][2] scratch: each query's 8 smallest candidates, packed
};
// ORBmatcher::SearchByProjection over already-projected queries (k_gridmatch.cu); frame f reads keys + f * key_stride,
// queries + f * q_stride; n / nq == nullptr: every frame has n_fixed / nq_fixed entries
struct ProjSearchArgs {
    const gforb_keypoint* keys; const uint8_t* desc; const float* u_right; const uint8_t* occupied; const int32_t* n;
    const gforb_proj_query* q; const uint8_t* qdesc; const int32_t* nq;
    int n_fixed, key_stride, nq_fixed, q_stride;
    const int32_t* cell_start; const int32_t* cell_idx;
    float min_x, min_y, w_inv, h_inv;
    float ratio;
    int th_dist, use_ratio, check_ori;
    float inv_sig2[32];          // mvInvLevelSigma2 for GFORB_PQ_REPROJ_GATE (ORBmatcher::Fuse)
    uint4* spec;                 // [frame][q_stride][2] scratch: each query's 8 smallest candidates, packed
    int32_t* key_query;          // [frame][key_stride]
    int32_t* key_dist;           // [frame][key_stride]
    int32_t* query_key;          // [frame][q_stride]
    int32_t* nmatches;           // [frame]
};
int launch_proj_search(const ProjSearchArgs& a, const float* bounds, int nframes, cudaStream_t s);   // 2 launches; -1: smem opt-in failed
// ORBmatcher::SearchByBoW over flattened DBoW2 feature vectors (k_gridmatch.cu); all pointers device
struct BowSearchArgs {
    const gforb_keypoint* k1; const uint8_t* d1; const uint8_t* valid1; int n1;
    const int32_t* node_id1; const int32_t* node_start1; const int32_t* node_idx1; int nn1;
    const gforb_keypoint* k2; const uint8_t* d2; const uint8_t* valid2; int n2;
    const int32_t* node_id2; const int32_t* node_start2; const int32_t* node_idx2; int nn2;
    int th_dist; float ratio; int check_ori;
    // SearchForTriangulation only (tri != 0): stereo flags, F12 row-major, epipole, pKF2's scale factors and sigma^2
    int tri; const uint8_t* stereo1; const uint8_t* stereo2;
    float F12[9]; float ex, ey; float sf[32]; float sig2[32];
    uint8_t* matched2;           // [n2] scratch
    int32_t* matches12;          // [n1]
    int32_t* nmatches;           // [1]
};
void launch_bow_search(const BowSearchArgs& a, cudaStream_t s);   // 2 memsets + 2 launches
// SearchBySim3 agreement test over the two searches' query_key arrays (k_gridmatch.cu); 1 memset + 1 launch
void launch_sim3_agree(const int32_t* match1, int n1, const int32_t* match2, int n2, int32_t* match12, int32_t* nfound,
                       cudaStream_t s);
int gridmatch_max_keys();   // 8192 keys per frame
int gridmatch_cells();      // 64 * 48
// bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}; counts == nullptr: every frame has n_fixed keys.  1 launch.
void launch_grid_build(const gforb_keypoint* keys, const int32_t* counts, int n_fixed, int key_stride, int nframes,
                       const float* bounds, int32_t* cell_start, int32_t* cell_idx, cudaStream_t s);
// queries: nq x {float x, y, r; int32 min_level, max_level}; out_start == nullptr counts, else fills.  1 launch.
void launch_area_query(const gforb_keypoint* keys, const float* bounds, const int32_t* cell_start, const int32_t* cell_idx,
                       const void* queries, int nq, int32_t* out_count, const int32_t* out_start, int32_t* out_idx,
                       cudaStream_t s);
int launch_init_search(const InitSearchArgs& a, const float* bounds, int npairs, cudaStream_t s);   // 2 launches; -1: smem opt-in failed
// input pre-processing (k_preprocess.cu): [16 -> 8 bit] -> remap -> gray; 1 launch.  fmt: gforb.h GFORB_PIX_*
int preprocess_bytes_per_pixel(int fmt);
void launch_preprocess(const uint8_t* src, int sw, int sh, size_t spitch, size_t sstride, int fmt, const float* mapx,
                       const float* mapy, uint8_t* dst, int dw, int dh, size_t dpitch, size_t dstride, int nimg, cudaStream_t s);
// keypoint undistortion (k_undistort.cu): cv::undistortPoints in double; K[9], dist[8], R[9], P[9] row-major.  1 launch each.
void launch_undistort_keys(const double* K, const double* dist, const double* R, const double* P, const gforb_keypoint* in,
                           const int32_t* counts, int n_fixed, int stride, int nimg, gforb_keypoint* out, cudaStream_t s);
void launch_undistort_xy(const double* K, const double* dist, const double* R, const double* P, const float* in, int n,
                         float* out, cudaStream_t s);
// MapPoint::ComputeDistinctiveDescriptors (k_match.cu): 1 launch; -1 = shared-memory opt-in failed
int distinctive_max_set();
int launch_distinctive(const uint8_t* desc, const int32_t* set_start, int npoints, int max_n, int32_t* best_idx,
                       int32_t* best_median, cudaStream_t s);
// multi-index hashing (k_mih.cu): substring keys, batched insertion (<= mih_max_batch() entries per launch), queries
int mih_max_batch();
void launch_mih_hash(const uint8_t* desc, int n, int T, int bits, uint32_t* keys, cudaStream_t s);
int launch_mih_insert(const uint8_t* desc, const int32_t* entry, int n, int T, int bits, int cap, int32_t* head, int32_t* pool_id,
                      int32_t* pool_next, int32_t* fill, uint32_t* hashed, uint8_t* success, int32_t* err, cudaStream_t s);
void launch_mih_query(const uint8_t* qdesc, int m, int T, int bits, int cap, const int32_t* head, const int32_t* pool_id,
                      const int32_t* pool_next, const int* take, int keep, int32_t* cand, int32_t* count, int32_t* addi,
                      int32_t* addi_count, int stride, cudaStream_t s);
size_t octree_smem_bytes(const Geom& g);
int fast_configure(const Geom& g);     // opt in to the FAST kernel's dynamic smem; returns cudaError_t as int
int octree_configure(const Geom& g);   // opt in to large dynamic smem; returns cudaError_t as int

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device and is a SET, not a raise: the opted-in size is kept
// as a high-water mark per (kernel, device) -- `have` is a zero-initialised static of the call site -- so that a second
// engine with a smaller footprint (another nfeatures, another image size) never lowers it under a live engine's
// launches.  The table is shared by every handle and handles are used from concurrent host threads (the reference's
// two-thread fork/join, src/Frame.cc:84-87): check-and-raise runs under one mutex.
inline std::mutex& smem_opt_in_mutex() {
    static std::mutex m;
    return m;
}
template <class K>
inline int opt_in_smem(K kernel, size_t smem, size_t (&have)[64]) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    std::lock_guard<std::mutex> lock(smem_opt_in_mutex());
    if (smem <= have[dev]) return 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return -1;
    have[dev] = smem;
    return 0;
}

void upload_constants();   // pattern + umax to __constant__ (once per process/device)

}  // namespace gforb
