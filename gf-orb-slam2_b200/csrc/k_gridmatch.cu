// Frame feature grid + windowed initialisation matcher.
//
// Reference:
//   Frame::AssignFeaturesToGrid / PosInGrid   src/Frame.cc:461-476, 645-656   (64 x 48 cells, include/Frame.h:92-93)
//   Frame::GetFeaturesInArea                  src/Frame.cc:593-643
//   ORBmatcher::SearchForInitialization       src/ORBmatcher.cc:520-633
//   ORBmatcher::ComputeThreeMaxima            src/ORBmatcher.cc:1723-1764
//
// Grid layout in HBM: the reference keeps mGrid[ix][iy] as 3,072 std::vector<size_t>, filled by push_back in
// ascending key index.  Here it is one CSR: `cell_idx` = key indices sorted by (ix, iy, index) and `cell_start`
// (3,073 offsets, cell id = ix * 48 + iy).  GetFeaturesInArea walks ix ascending, then iy ascending, then the cell
// in push_back order -- which is exactly ascending CSR position, and for one ix the cells iy0..iy1 are ONE
// contiguous CSR run.  So "candidate order" (it decides which of two equal Hamming minima wins) is the CSR position
// and needs no per-query list, no ballot and no atomics.
//
// The searches (SearchForInitialization, SearchByProjection, SearchByBoW, SearchForTriangulation) are described at their
// sections below; the ones with a sequential rule in the reference share one scheme: a grid-wide scan kernel keeps each
// query's 8 smallest candidates, a CTA per frame replays the queries in order against claims held in shared memory.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kGCols = 64, kGRows = 48, kGCells = kGCols * kGRows;
constexpr int kHisto = 30;                      // HISTO_LENGTH, src/ORBmatcher.cc:39
constexpr int kPosBits = 13;                    // CSR position bits inside a key: n <= 8192
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr int kGridThreads = 256, kSearchThreads = 256;

struct GridView {
    const int32_t* start;      // [kGCells + 1]
    const int32_t* idx;        // [n] key indices in (ix, iy, index) order
    float min_x, min_y, w_inv, h_inv;
};

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return (__popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z)) +
           (__popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y)) +
           (__popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w));
}

// cell window of GetFeaturesInArea (src/Frame.cc:598-612); false = the reference returns an empty vector
__device__ __forceinline__ bool area_window(const GridView& g, float x, float y, float r, int& x0, int& x1, int& y0, int& y1) {
    x0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, g.min_x), r), g.w_inv)));
    if (x0 >= kGCols) return false;
    x1 = min(kGCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, g.min_x), r), g.w_inv)));
    if (x1 < 0) return false;
    y0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, g.min_y), r), g.h_inv)));
    if (y0 >= kGRows) return false;
    y1 = min(kGRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, g.min_y), r), g.h_inv)));
    return y1 >= 0;
}

// level + square-window test of one grid entry (src/Frame.cc:623-637)
__device__ __forceinline__ bool area_accepts(const gforb_keypoint* __restrict__ keys, int i, float x, float y, float r,
                                             int lo, int hi) {
    const int oct = __ldg(&keys[i].octave);
    if (lo > 0 || hi >= 0) {
        if (oct < lo) return false;
        if (hi >= 0 && oct > hi) return false;
    }
    const float dx = __fsub_rn(__ldg(&keys[i].x), x), dy = __fsub_rn(__ldg(&keys[i].y), y);
    return fabsf(dx) < r && fabsf(dy) < r;
}
}  // namespace

// ---------------------------------------------------------------------------------------------------- grid build
// One CTA per frame, a counting sort by cell: histogram of cell ids (shared atomics), exclusive scan = cell_start,
// scatter into the cells in any order, then every entry finds its place inside its cell by counting the smaller key
// indices there (cells hold a handful of keys, so this is a few compares; the count is what "push_back in ascending i",
// src/Frame.cc:468-475, produces).  Keys outside the grid (PosInGrid false, :652-653) are dropped.
// dynamic shared: cell_of[n] u16 | slot[n] u16
__global__ void __launch_bounds__(kGridThreads)
grid_build_kernel(const gforb_keypoint* __restrict__ keys, const int32_t* __restrict__ counts, int n_fixed, int key_stride,
                  float min_x, float min_y, float w_inv, float h_inv, int32_t* __restrict__ cell_start,
                  int32_t* __restrict__ cell_idx, int ncap) {
    extern __shared__ uint16_t gsm[];
    __shared__ int start[kGCells + 1];
    __shared__ int fill[kGCells];
    __shared__ int warp_tot[kGridThreads / 32];
    uint16_t* cell_of = gsm;
    uint16_t* slot = gsm + ncap;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n = counts ? min(counts[f], key_stride) : n_fixed;
    const gforb_keypoint* k = keys + (size_t)f * key_stride;
    int32_t* gstart = cell_start + (size_t)f * (kGCells + 1);
    int32_t* gidx = cell_idx + (size_t)f * key_stride;
    for (int c = tid; c < kGCells; c += kGridThreads) fill[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {
        const int px = (int)roundf(__fmul_rn(__fsub_rn(__ldg(&k[i].x), min_x), w_inv));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(__ldg(&k[i].y), min_y), h_inv));
        const bool in = px >= 0 && px < kGCols && py >= 0 && py < kGRows;
        cell_of[i] = in ? (uint16_t)(px * kGRows + py) : (uint16_t)0xFFFF;
        if (in) atomicAdd(&fill[px * kGRows + py], 1);
    }
    __syncthreads();
    // exclusive scan of the 3,072 counts: 12 consecutive cells per thread, warp scan, scan of the 8 warp totals
    constexpr int kPer = kGCells / kGridThreads;
    int local = 0;
#pragma unroll
    for (int j = 0; j < kPer; ++j) local += fill[tid * kPer + j];
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    int base = incl - local;
    for (int w = 0; w < wid; ++w) base += warp_tot[w];
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int c = tid * kPer + j, cnt = fill[c];
        start[c] = base;
        gstart[c] = base;
        fill[c] = base;                                   // running fill pointer of the cell
        base += cnt;
    }
    if (tid == kGridThreads - 1) { start[kGCells] = base; gstart[kGCells] = base; }    // keys inside the grid
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {
        const int c = cell_of[i];
        if (c != 0xFFFF) slot[atomicAdd(&fill[c], 1)] = (uint16_t)i;
    }
    __syncthreads();
    const int n_in = start[kGCells];
    for (int p = tid; p < n_in; p += kGridThreads) {
        const int e = slot[p], c = cell_of[e];
        const int s0 = start[c], s1 = start[c + 1];
        int r = 0;
        for (int j = s0; j < s1; ++j) r += slot[j] < e;
        gidx[s0 + r] = e;
    }
}

// ---------------------------------------------------------------------------------------- GetFeaturesInArea batch
struct AreaQuery { float x, y, r; int32_t min_level, max_level; };

// warp per query.  out_start == nullptr: write the count to out_count[q]; else write the indices at out_start[q].
__global__ void __launch_bounds__(256)
area_query_kernel(const gforb_keypoint* __restrict__ keys, GridView g, const AreaQuery* __restrict__ qs, int nq,
                  int32_t* __restrict__ out_count, const int32_t* __restrict__ out_start, int32_t* __restrict__ out_idx) {
    const int lane = threadIdx.x & 31;
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= nq) return;
    const AreaQuery a = qs[q];
    int x0, x1, y0, y1, total = 0;
    const int base = out_start ? out_start[q] : 0;
    if (area_window(g, a.x, a.y, a.r, x0, x1, y0, y1)) {
        for (int ix = x0; ix <= x1; ++ix) {
            const int j0 = __ldg(g.start + ix * kGRows + y0), j1 = __ldg(g.start + ix * kGRows + y1 + 1);
            for (int jb = j0; jb < j1; jb += 32) {                       // warp-uniform trip count (ballot inside)
                const int j = jb + lane;
                int i = -1;
                bool ok = false;
                if (j < j1) { i = __ldg(g.idx + j); ok = area_accepts(keys, i, a.x, a.y, a.r, a.min_level, a.max_level); }
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, ok);
                if (ok && out_start) out_idx[base + total + __popc(m & ((1u << lane) - 1))] = i;
                total += __popc(m);
            }
        }
    }
    if (!out_start && lane == 0) out_count[q] = total;
}

// ------------------------------------------------------------------------------------------ SearchByProjection
// src/ORBmatcher.cc:155-241 (map points), :1440-1593 (last frame), :1595-1721 (key frame), folded as include/gforb.h
// describes.  Sequential rule in the reference: a key taken by a query whose map point has observations is skipped by
// every later query.  Parallel restatement in two kernels:
//   proj_scan_kernel     warp per query, whole grid: the query's kTopK smallest candidates (distance, then window
//                        order) against the INITIAL occupancy, each packed as dist << 19 | CSR position << 6 | level;
//   proj_resolve_kernel  CTA per frame, lists + claims + flags in shared memory, claims indexed by CSR POSITION so the
//                        replay chases no global pointer.  The CTA replays the queries in order, W per step (thread =
//                        query): a thread walks its list past the positions already claimed; its answer is the
//                        sequential one unless an earlier thread of the same step takes its best (or, with the ratio
//                        test, second-best) position (shared atomicMin of the thread number per position).  The
//                        threads in front of the first such thread commit together and the next step starts AT that
//                        thread, which as thread 0 sees every claim.  W = 256 while steps get >= 32 queries through,
//                        else 32.  When the query at the head has used its list up, all 8 warps refresh the lists of
//                        EVERY not-yet-replayed query in that state (claims only grow, so a list made under the
//                        current claims stays valid), then the replay resumes.  What a query took stays in slot 0 of
//                        its list; key indices, holders (atomicMax = last taker), scores, rotation bins, histogram,
//                        three maxima and removal are written afterwards by all threads.
// Measured (B200, 1,008 keys x 1,008 queries, last-frame form): 256 frames per launch 0.24-0.27 ms = 0.96-1.07 M frames/s;
// one frame 48-83 us (grid 7, scan 10, resolve 30-65) against 0.55 ms for the sequential CPU oracle.
namespace {
constexpr int kTopK = 8;
constexpr int kLvlBits = 6;                                     // levels < 64
constexpr int kDistShift = kPosBits + kLvlBits;
constexpr uint32_t kTrunc = 0xFFFFFFFEu;                        // "more candidates may follow": list cut here

__device__ __forceinline__ int key_pos(uint32_t e) { return (int)((e >> kLvlBits) & ((1u << kPosBits) - 1)); }

// kByPos: `blocked` is indexed by CSR position (the replay's claims), else by key index (the caller's occupancy)
template <bool kByPos>
__device__ __forceinline__ void proj_scan(const GridView& g, const gforb_keypoint* __restrict__ keys,
                                          const uint4* __restrict__ desc, const float* __restrict__ u_right,
                                          const uint8_t* blocked, const gforb_proj_query& q, const uint4& a0,
                                          const uint4& a1, int lane, const float* inv_sig2, uint32_t (&top)[kTopK],
                                          const uint16_t* cdist = nullptr, bool keep_all = false) {
    uint32_t best = kNone, second = kNone;
    int cnt = 0;
    const bool gate = (q.flags & GFORB_PQ_RIGHT_GATE) && u_right;
    int x0, x1, y0, y1;
    if (area_window(g, q.x, q.y, q.r, x0, x1, y0, y1)) {
        // The window's columns are CSR runs [j0, j0 + len); all of them (<= 64) are fetched at once, two per lane, and
        // the entries of all runs are dealt round the lanes as ONE sequence t = 0 .. total-1 (column by binary search
        // in the prefix sums, via shuffles), so the warp walks ceil(total / 32) steps instead of one step per column.
        const int ncol = x1 - x0 + 1;
        int j0a = 0, lena = 0, j0b = 0, lenb = 0;
        if (lane < ncol) {
            j0a = __ldg(g.start + (x0 + lane) * kGRows + y0);
            lena = __ldg(g.start + (x0 + lane) * kGRows + y1 + 1) - j0a;
        }
        if (lane + 32 < ncol) {
            j0b = __ldg(g.start + (x0 + lane + 32) * kGRows + y0);
            lenb = __ldg(g.start + (x0 + lane + 32) * kGRows + y1 + 1) - j0b;
        }
        int pa = lena, pb = lenb;                     // inclusive prefix sums of the run lengths
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int va = __shfl_up_sync(0xFFFFFFFFu, pa, o), vb = __shfl_up_sync(0xFFFFFFFFu, pb, o);
            if (lane >= o) { pa += va; pb += vb; }
        }
        const int tot_a = __shfl_sync(0xFFFFFFFFu, pa, 31);
        pb += tot_a;
        const int total = __shfl_sync(0xFFFFFFFFu, pb, 31);
        for (int tb = 0; tb < total; tb += 32) {
            const int t = tb + lane;
            int ca = 0, cb = 0;                       // first column whose inclusive prefix exceeds t, in each half
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const int va = __shfl_sync(0xFFFFFFFFu, pa, ca + step - 1), vb = __shfl_sync(0xFFFFFFFFu, pb, cb + step - 1);
                if (va <= t) ca += step;
                if (vb <= t) cb += step;
            }
            const bool in_a = t < tot_a;
            const int src = in_a ? ca : min(cb, 31);
            const int ra = __shfl_sync(0xFFFFFFFFu, j0a + lena - pa, src);          // j0 - exclusive prefix of the run
            const int rb = __shfl_sync(0xFFFFFFFFu, j0b + lenb - pb, src);
            if (t >= total) continue;
            const int j = (in_a ? ra : rb) + t;
            const int i = __ldg(g.idx + j);
            if (!area_accepts(keys, i, q.x, q.y, q.r, q.min_level, q.max_level)) continue;
            if (blocked && blocked[kByPos ? j : i]) continue;                      // :197-199, :1519-1521, :1662
            if (gate) {                                                            // :201-206, :1523-1529
                const float ur = __ldg(u_right + i);
                if (ur > 0.0f && fabsf(__fsub_rn(q.xr, ur)) > q.r) continue;
            }
            if (q.flags & GFORB_PQ_REPROJ_GATE) {                                  // ORBmatcher::Fuse, :1022-1046
                const float ex = __fsub_rn(q.x, __ldg(&keys[i].x)), ey = __fsub_rn(q.y, __ldg(&keys[i].y));
                const float inv = inv_sig2[__ldg(&keys[i].octave) & 31];
                const float ur = u_right ? __ldg(u_right + i) : -1.0f;
                float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                if (ur >= 0.0f) {
                    const float er = __fsub_rn(q.xr, ur);
                    e2 = __fadd_rn(e2, __fmul_rn(er, er));
                    if ((double)__fmul_rn(e2, inv) > 7.8) continue;
                } else if ((double)__fmul_rn(e2, inv) > 5.99) continue;
            }
            const int dist = hamming256(a0, a1, __ldg(desc + 2 * (size_t)i), __ldg(desc + 2 * (size_t)i + 1));
            if (dist >= 256 && !keep_all) continue;                                // never below the initial 256
            if (cdist && (int)cdist[j] <= dist) continue;                          // vMatchedDistance[i2] <= dist, :559-560
            const uint32_t key = ((uint32_t)dist << kDistShift) | ((uint32_t)j << kLvlBits) |
                                 ((uint32_t)__ldg(&keys[i].octave) & ((1u << kLvlBits) - 1));
            ++cnt;
            second = min(second, max(best, key));
            best = min(best, key);
        }
    }
    // the warp's kTopK smallest, popped from the lanes' two smallest; a lane that gave both and had seen more may hide
    // the next one: the list is cut there
    bool trunc = false;
#pragma unroll
    for (int k = 0; k < kTopK; ++k) {
        uint32_t m = __reduce_min_sync(0xFFFFFFFFu, best);
        if (trunc) m = kTrunc;
        top[k] = m;
        bool spent = false;
        if (m < kTrunc && best == m) { best = second; second = kNone; spent = best == kNone && cnt > 2; }
        trunc = trunc || __any_sync(0xFFFFFFFFu, spent);
    }
}

__device__ __forceinline__ GridView proj_grid(const ProjSearchArgs& a, int f) {
    return GridView{a.cell_start + (size_t)f * (kGCells + 1), a.cell_idx + (size_t)f * a.key_stride, a.min_x, a.min_y,
                    a.w_inv, a.h_inv};
}
}  // namespace

__global__ void __launch_bounds__(256)
proj_scan_kernel(const ProjSearchArgs a) {
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int qi = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nq = a.nq ? min(a.nq[f], a.q_stride) : a.nq_fixed;
    if (qi >= nq) return;
    const int n = a.n ? min(a.n[f], a.key_stride) : a.n_fixed;
    const gforb_proj_query q = a.q[(size_t)f * a.q_stride + qi];
    uint32_t top[kTopK];
#pragma unroll
    for (int k = 0; k < kTopK; ++k) top[k] = kNone;
    if (n > 0 && (q.flags & GFORB_PQ_ACTIVE)) {
        const uint4* qd = reinterpret_cast<const uint4*>(a.qdesc) + 2 * ((size_t)f * a.q_stride + qi);
        const uint4 a0 = __ldg(qd), a1 = __ldg(qd + 1);
        proj_scan<false>(proj_grid(a, f), a.keys + (size_t)f * a.key_stride,
                         reinterpret_cast<const uint4*>(a.desc) + 2 * (size_t)f * a.key_stride,
                         a.u_right ? a.u_right + (size_t)f * a.key_stride : nullptr,
                         a.occupied ? a.occupied + (size_t)f * a.key_stride : nullptr, q, a0, a1, lane, a.inv_sig2, top);
    }
    if (lane == 0) {
        uint4* o = a.spec + 2 * ((size_t)f * a.q_stride + qi);
        o[0] = make_uint4(top[0], top[1], top[2], top[3]);
        o[1] = make_uint4(top[4], top[5], top[6], top[7]);
    }
}

// dynamic shared: lists[topk][nq] u32 (room for qcap queries) | tent[8192] u32
__global__ void __launch_bounds__(kSearchThreads)
proj_resolve_kernel(const ProjSearchArgs a, int topk, int qcap) {
    extern __shared__ uint32_t lists[];
    uint32_t* tent = lists + (size_t)topk * qcap;               // by CSR position: first lane about to take it this step
    __shared__ uint8_t claimed[1 << kPosBits];                  // by CSR position
    __shared__ uint8_t bin[1 << kPosBits];
    __shared__ uint8_t qflag[1 << kPosBits];
    __shared__ int hist[kHisto];
    __shared__ int keep_bin[3];
    __shared__ int total, s_base, s_done, s_first, s_win;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n = a.n ? min(a.n[f], a.key_stride) : a.n_fixed;
    const int nq = a.nq ? min(a.nq[f], a.q_stride) : a.nq_fixed;
    const GridView g = proj_grid(a, f);
    const gforb_keypoint* keys = a.keys + (size_t)f * a.key_stride;
    const uint8_t* occ = a.occupied ? a.occupied + (size_t)f * a.key_stride : nullptr;
    const gforb_proj_query* qs = a.q + (size_t)f * a.q_stride;
    const uint32_t* spec = reinterpret_cast<const uint32_t*>(a.spec + 2 * (size_t)f * a.q_stride);
    int32_t* key_query = a.key_query + (size_t)f * a.key_stride;
    int32_t* key_dist = a.key_dist + (size_t)f * a.key_stride;
    int32_t* query_key = a.query_key + (size_t)f * a.q_stride;

    const int n_in = n > 0 ? __ldg(g.start + kGCells) : 0;      // keys inside the grid = CSR length
    for (int j = tid; j < n_in; j += kSearchThreads) { claimed[j] = occ ? (occ[__ldg(g.idx + j)] != 0) : 0; tent[j] = kNone; }
    for (int i = tid; i < n; i += kSearchThreads) { key_query[i] = -1; key_dist[i] = 0; }
    for (int i = tid; i < nq; i += kSearchThreads) { bin[i] = 0xFF; query_key[i] = -1; qflag[i] = (uint8_t)qs[i].flags; }
    for (int t = tid; t < nq * topk; t += kSearchThreads) {      // lists[k][query]: lanes = queries, no bank conflicts
        const int qi = t / topk, k = t - qi * topk;
        lists[k * nq + qi] = spec[qi * kTopK + k];
    }
    if (tid < kHisto) hist[tid] = 0;
    if (tid == 0) { total = 0; s_base = 0; s_done = 0; s_first = kSearchThreads; s_win = 32; }
    __syncthreads();

    const uint32_t full = 0xFFFFFFFFu;
    // threshold (:228, :1532, :1674) and the same-level ratio test (:230) on packed entries
    auto accepts = [&](uint32_t b, uint32_t s2) {
        if (b >= kTrunc) return false;
        const int bd = (int)(b >> kDistShift);
        if (bd > a.th_dist) return false;
        if (!a.use_ratio || s2 >= kTrunc) return true;                              // bestLevel2 = -1
        const bool same_level = ((b ^ s2) & ((1u << kLvlBits) - 1)) == 0;
        return !(same_level && (float)bd > __fmul_rn(a.ratio, (float)(s2 >> kDistShift)));
    };
    // first (and, with the ratio test, second) entry of query qi's list whose position is still free;
    // true = the list ran out before that answer was complete.  All entries and then all their claim flags are
    // fetched at once (two shared-memory latencies), the selection itself runs in registers.
    auto walk = [&](int qi, uint32_t& b, uint32_t& s2) {
        uint32_t e[kTopK];
        bool cl[kTopK];
#pragma unroll
        for (int k = 0; k < kTopK; ++k) e[k] = k < topk ? lists[k * nq + qi] : kTrunc;
#pragma unroll
        for (int k = 0; k < kTopK; ++k) cl[k] = e[k] < kTrunc && claimed[key_pos(e[k])];
        b = kNone; s2 = kNone;
        bool need = false, done = false;
#pragma unroll
        for (int k = 0; k < kTopK; ++k) {
            if (done) continue;
            if (e[k] == kNone) done = true;
            else if (e[k] == kTrunc) { need = true; done = true; }
            else if (!cl[k]) {
                if (b == kNone) { b = e[k]; done = !a.use_ratio; }
                else { s2 = e[k]; done = true; }
            }
        }
        return need || !done;
    };
    // Replay, the whole CTA per step.  The step covers W queries: 256 while steps get at least 32 queries through, 32 (one
    // warp's worth of walking; the other threads only keep the barriers) while they do not, so that dense competition
    // does not multiply the walking by eight.
    while (true) {
        const int base = s_base, W = s_win;
        if (base >= nq) break;
        const int qi = base + tid;
        const bool valid = tid < W && qi < nq;
        uint32_t b = kNone, s2 = kNone;
        const bool need = valid && walk(qi, b, s2);
        if (tid == 0) s_done = need;                                                    // query `base` needs a fresh list
        const bool acc = valid && !need && accepts(b, s2);
        const bool blocks = acc && (qflag[valid ? qi : 0] & GFORB_PQ_BLOCKS);
        const int bpos = b < kTrunc ? key_pos(b) : -2, spos = (a.use_ratio && s2 < kTrunc) ? key_pos(s2) : -2;
        // does an earlier thread of this step take my best or second-best position?  Each position about to be taken
        // gets the number of its first taker in tent[]; everyone looks up its own two
        if (blocks) atomicMin(&tent[bpos], (uint32_t)tid);
        __syncthreads();
        if (s_done) {
            // refresh round, all warps: every query not yet replayed whose list is used up under the claims as they
            // stand gets a new list (claims only grow, so a list made now stays a valid list later)
            if (blocks) tent[bpos] = kNone;
            for (int q0 = base + wid * 32; q0 < nq; q0 += kSearchThreads) {
                const int qj = q0 + lane;
                uint32_t tb, ts;
                uint32_t m = __ballot_sync(full, qj < nq && walk(qj, tb, ts));
                while (m) {
                    const int qr = q0 + __ffs(m) - 1;
                    m &= m - 1;
                    const gforb_proj_query q = qs[qr];
                    const uint4* qd = reinterpret_cast<const uint4*>(a.qdesc) + 2 * ((size_t)f * a.q_stride + qr);
                    const uint4 a0 = __ldg(qd), a1 = __ldg(qd + 1);
                    uint32_t top[kTopK];
                    proj_scan<true>(g, keys, reinterpret_cast<const uint4*>(a.desc) + 2 * (size_t)f * a.key_stride,
                                    a.u_right ? a.u_right + (size_t)f * a.key_stride : nullptr, claimed, q, a0, a1, lane, a.inv_sig2, top);
#pragma unroll
                    for (int k = 0; k < kTopK; ++k)
                        if (lane == 0 && k < topk) lists[k * nq + qr] = top[k];
                }
            }
            __syncthreads();
            continue;                                                                   // same base, fresh lists
        }
        const bool bad = need || (bpos >= 0 && tent[bpos] < (uint32_t)tid) || (spos >= 0 && tent[spos] < (uint32_t)tid);
        if (bad) atomicMin(&s_first, tid);                                              // thread 0 is never bad here
        __syncthreads();
        const int nb = min(s_first, W);
        if (blocks) tent[bpos] = kNone;
        // slot 0 of a replayed query's list now holds what it took (kNone: nothing); the global side of the
        // assignment (key index, holder, score, rotation bin) is written after the replay, by all threads
        if (valid && tid < nb) {
            lists[qi] = acc ? b : kNone;
            if (blocks) claimed[bpos] = 1;
        }
        __syncthreads();
        if (tid == 0) {
            s_base = base + nb;
            s_first = kSearchThreads;
            s_win = nb >= 32 ? kSearchThreads : 32;                 // wide while a step gets at least a warp's worth through
        }
        __syncthreads();
    }

    // the assignments (:232-236, :1534-1563): key index of each taken position, rotation bin, and the key's holder =
    // the LAST query that took it (an earlier taker that did not block is overwritten, as in the reference)
    for (int q = tid; q < nq; q += kSearchThreads) {
        const uint32_t e = lists[q];
        if (e >= kTrunc) continue;
        const int ib = __ldg(g.idx + key_pos(e));
        query_key[q] = ib;
        atomicMax(&key_query[ib], q);
        if (a.check_ori) {                                                   // :1553-1562, :1681-1690
            float rot = __fsub_rn(qs[q].angle, __ldg(&keys[ib].angle));
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bb = (int)roundf(__fmul_rn(rot, 1.0f / kHisto));
            bb = bb == kHisto ? 0 : bb;
            bin[q] = (uint8_t)bb;
            atomicAdd(&hist[bb], 1);
        }
    }
    __syncthreads();
    for (int q = tid; q < nq; q += kSearchThreads) {
        const uint32_t e = lists[q];
        if (e >= kTrunc) continue;
        const int ib = query_key[q];
        if (key_query[ib] == q) key_dist[ib] = (int)(e >> kDistShift);       // mvpMatchScore, :235
    }
    if (a.check_ori) {                                                       // :1568-1590, :1696-1718
        if (tid == 0) {                                                      // ComputeThreeMaxima, :1723-1764
            int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < kHisto; ++i) {
                const int s = hist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
                else if (s > max3) { max3 = s; i3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { i3 = -1; }
            keep_bin[0] = i1; keep_bin[1] = i2; keep_bin[2] = i3;
        }
    }
    __syncthreads();
    int mine = 0;
    for (int q = tid; q < nq; q += kSearchThreads) {
        int k = query_key[q];
        if (k >= 0 && a.check_ori) {
            const int b = bin[q];
            if (b != keep_bin[0] && b != keep_bin[1] && b != keep_bin[2]) { key_query[k] = -2; query_key[q] = -1; k = -1; }
        }
        mine += k >= 0;
    }
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if (lane == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (tid == 0) a.nmatches[f] = total;
}

// ------------------------------------------------------------------------------------ SearchForInitialization
// src/ORBmatcher.cc:520-633.  Sequential in the reference: query i1 ignores a target i2 that an earlier query holds
// with a distance <= its own (vMatchedDistance, :559-560) and may take a target over (:578-582).  Device restatement
// with the machinery of SearchByProjection above:
//   init_scan_kernel     warp per level-0 query, whole grid: its 8 smallest candidates (window vbPrevMatched +- windowSize,
//                        level 0) against an empty claim table;
//   init_resolve_kernel  CTA per frame pair, in-order replay W queries per step.  A claim holds the claimed DISTANCE
//                        (cdist[position], 0xFFFF = free): a candidate is visible while its claim is larger than its
//                        own distance, so a thread's best / second-best are the first two visible entries of its list.
//                        Claims only get smaller, hence hidden entries stay hidden and only a claim on the thread's
//                        best or second-best position by an earlier thread of the step invalidates its answer.
//                        holder[position] = last taker (take-over), bins of every accepted query (rotHist counts
//                        matches that were later taken over, as push_back does), three maxima, removal, vbPrevMatched
//                        update and the count by all threads.
__global__ void __launch_bounds__(256)
init_scan_kernel(const InitSearchArgs a) {
    const int p = blockIdx.y, lane = threadIdx.x & 31, q = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int n1 = a.n1 ? min(a.n1[p], a.stride1) : a.n1_fixed;
    if (q >= n1) return;
    const int n2 = a.n2 ? min(a.n2[p], a.stride2) : a.n2_fixed;
    const gforb_keypoint* k1 = a.k1 + (size_t)p * a.stride1;
    uint32_t top[kTopK];
#pragma unroll
    for (int k = 0; k < kTopK; ++k) top[k] = kNone;
    if (n2 > 0 && __ldg(&k1[q].octave) <= 0) {                            // level1 > 0: continue (:537-538)
        const uint4* d1 = reinterpret_cast<const uint4*>(a.d1) + 2 * (size_t)p * a.stride1;
        const float* prev = a.prev_xy + 2 * (size_t)p * a.stride1;
        const GridView g{a.cell_start + (size_t)p * (kGCells + 1), a.cell_idx + (size_t)p * a.stride2, a.min_x, a.min_y, a.w_inv, a.h_inv};
        const gforb_proj_query pq{prev[2 * q], prev[2 * q + 1], a.window, 0.f, 0.f, 0, 0, GFORB_PQ_ACTIVE};   // level1, level1 = 0, 0
        const uint4 a0 = __ldg(d1 + 2 * (size_t)q), a1 = __ldg(d1 + 2 * (size_t)q + 1);
        proj_scan<false>(g, a.k2 + (size_t)p * a.stride2, reinterpret_cast<const uint4*>(a.d2) + 2 * (size_t)p * a.stride2, nullptr,
                         nullptr, pq, a0, a1, lane, nullptr, top, nullptr, true);
    }
    if (lane == 0) {
        uint4* o = reinterpret_cast<uint4*>(a.spec) + 2 * ((size_t)p * a.stride1 + q);
        o[0] = make_uint4(top[0], top[1], top[2], top[3]);
        o[1] = make_uint4(top[4], top[5], top[6], top[7]);
    }
}

// dynamic shared: lists[topk][n1] u32 (room for qcap queries) | tent[8192] u32
__global__ void __launch_bounds__(kSearchThreads)
init_resolve_kernel(const InitSearchArgs a, int topk, int qcap) {
    extern __shared__ uint32_t lists[];
    uint32_t* tent = lists + (size_t)topk * qcap;
    __shared__ uint16_t cdist[1 << kPosBits];                   // by CSR position: distance the position is held with
    __shared__ uint16_t holder[1 << kPosBits];                  // by CSR position: query holding it
    __shared__ uint8_t bin[1 << kPosBits];
    __shared__ int hist[kHisto];
    __shared__ int keep_bin[3];
    __shared__ int total, s_base, s_done, s_first, s_win;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n1 = a.n1 ? min(a.n1[p], a.stride1) : a.n1_fixed;
    const int n2 = a.n2 ? min(a.n2[p], a.stride2) : a.n2_fixed;
    const gforb_keypoint* k1 = a.k1 + (size_t)p * a.stride1;
    const gforb_keypoint* k2 = a.k2 + (size_t)p * a.stride2;
    const uint4* d1 = reinterpret_cast<const uint4*>(a.d1) + 2 * (size_t)p * a.stride1;
    const uint4* d2 = reinterpret_cast<const uint4*>(a.d2) + 2 * (size_t)p * a.stride2;
    float* prev = a.prev_xy + 2 * (size_t)p * a.stride1;
    int32_t* m12 = a.matches12 + (size_t)p * a.stride1;
    const GridView g{a.cell_start + (size_t)p * (kGCells + 1), a.cell_idx + (size_t)p * a.stride2, a.min_x, a.min_y, a.w_inv, a.h_inv};
    const uint32_t* spec = reinterpret_cast<const uint32_t*>(a.spec) + (size_t)kTopK * p * a.stride1;
    const int nq = n1;

    const int n_in = n2 > 0 ? __ldg(g.start + kGCells) : 0;
    for (int j = tid; j < n_in; j += kSearchThreads) { cdist[j] = 0xFFFF; tent[j] = kNone; }
    for (int i = tid; i < nq; i += kSearchThreads) { bin[i] = 0xFF; m12[i] = -1; }
    for (int t = tid; t < nq * topk; t += kSearchThreads) {
        const int qi = t / topk, k = t - qi * topk;
        lists[k * nq + qi] = spec[qi * kTopK + k];
    }
    if (tid < kHisto) hist[tid] = 0;
    if (tid == 0) { total = 0; s_base = 0; s_done = 0; s_first = kSearchThreads; s_win = 32; }
    __syncthreads();

    const uint32_t full = 0xFFFFFFFFu;
    // first and second entry of query qi's list still visible under the claims; true = the list ran out first
    auto walk = [&](int qi, uint32_t& b, uint32_t& s2) {
        uint32_t e[kTopK];
        bool hid[kTopK];
#pragma unroll
        for (int k = 0; k < kTopK; ++k) e[k] = k < topk ? lists[k * nq + qi] : kTrunc;
#pragma unroll
        for (int k = 0; k < kTopK; ++k) hid[k] = e[k] < kTrunc && (uint32_t)cdist[key_pos(e[k])] <= (e[k] >> kDistShift);   // :559-560
        b = kNone; s2 = kNone;
        bool need = false, done = false;
#pragma unroll
        for (int k = 0; k < kTopK; ++k) {
            if (done) continue;
            if (e[k] == kNone) done = true;
            else if (e[k] == kTrunc) { need = true; done = true; }
            else if (!hid[k]) {
                if (b == kNone) b = e[k];
                else { s2 = e[k]; done = true; }
            }
        }
        return need || !done;
    };
    // src/ORBmatcher.cc:574-576
    auto accepts = [&](uint32_t b, uint32_t s2) {
        if (b >= kTrunc) return false;
        const int bd = (int)(b >> kDistShift);
        const int sd = s2 < kTrunc ? (int)(s2 >> kDistShift) : INT32_MAX;
        return bd <= a.th_low && (float)bd < __fmul_rn((float)sd, a.ratio);
    };
    while (true) {
        const int base = s_base, W = s_win;
        if (base >= nq) break;
        const int qi = base + tid;
        const bool valid = tid < W && qi < nq;
        uint32_t b = kNone, s2 = kNone;
        const bool need = valid && walk(qi, b, s2);
        if (tid == 0) s_done = need;
        const bool acc = valid && !need && accepts(b, s2);
        const int bpos = b < kTrunc ? key_pos(b) : -2, spos = s2 < kTrunc ? key_pos(s2) : -2;
        if (acc) atomicMin(&tent[bpos], (uint32_t)tid);
        __syncthreads();
        if (s_done) {                                            // refresh round: new lists under the claims as they stand
            if (acc) tent[bpos] = kNone;
            for (int q0 = base + wid * 32; q0 < nq; q0 += kSearchThreads) {
                const int qj = q0 + lane;
                uint32_t tb, ts;
                uint32_t m = __ballot_sync(full, qj < nq && walk(qj, tb, ts));
                while (m) {
                    const int qr = q0 + __ffs(m) - 1;
                    m &= m - 1;
                    const gforb_proj_query pq{prev[2 * qr], prev[2 * qr + 1], a.window, 0.f, 0.f, 0, 0, GFORB_PQ_ACTIVE};
                    const uint4 a0 = __ldg(d1 + 2 * (size_t)qr), a1 = __ldg(d1 + 2 * (size_t)qr + 1);
                    uint32_t top[kTopK];
                    proj_scan<true>(g, k2, d2, nullptr, nullptr, pq, a0, a1, lane, nullptr, top, cdist, true);
#pragma unroll
                    for (int k = 0; k < kTopK; ++k)
                        if (lane == 0 && k < topk) lists[k * nq + qr] = top[k];
                }
            }
            __syncthreads();
            continue;
        }
        const bool bad = need || (bpos >= 0 && tent[bpos] < (uint32_t)tid) || (spos >= 0 && tent[spos] < (uint32_t)tid);
        if (bad) atomicMin(&s_first, tid);
        __syncthreads();
        const int nb = min(s_first, W);
        if (acc) tent[bpos] = kNone;
        if (valid && tid < nb) {
            lists[qi] = acc ? b : kNone;                         // slot 0: what the query took
            if (acc) { cdist[bpos] = (uint16_t)(b >> kDistShift); holder[bpos] = (uint16_t)qi; }      // :578-586
        }
        __syncthreads();
        if (tid == 0) {
            s_base = base + nb;
            s_first = kSearchThreads;
            s_win = nb >= 32 ? kSearchThreads : 32;
        }
        __syncthreads();
    }

    // matches of the current holders, rotation bins of every accepted query (:588-597)
    for (int q = tid; q < nq; q += kSearchThreads) {
        const uint32_t e = lists[q];
        if (e >= kTrunc) continue;
        const int pos = key_pos(e), ib = __ldg(g.idx + pos);
        if (holder[pos] == q) m12[q] = ib;
        if (a.check_ori) {
            float rot = __fsub_rn(__ldg(&k1[q].angle), __ldg(&k2[ib].angle));
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bb = (int)roundf(__fmul_rn(rot, 1.0f / kHisto));
            bb = bb == kHisto ? 0 : bb;
            bin[q] = (uint8_t)bb;
            atomicAdd(&hist[bb], 1);
        }
    }
    __syncthreads();
    if (a.check_ori) {                                                       // :601-622
        if (tid == 0) {                                                      // ComputeThreeMaxima, :1723-1764
            int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < kHisto; ++i) {
                const int s = hist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
                else if (s > max3) { max3 = s; i3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { i3 = -1; }
            keep_bin[0] = i1; keep_bin[1] = i2; keep_bin[2] = i3;
        }
        __syncthreads();
    }
    int mine = 0;
    for (int q = tid; q < nq; q += kSearchThreads) {                         // :625-627 and the count
        int m = m12[q];
        if (m >= 0 && a.check_ori) {
            const int bq = bin[q];
            if (bq != keep_bin[0] && bq != keep_bin[1] && bq != keep_bin[2]) { m = -1; m12[q] = -1; }
        }
        if (m >= 0) {
            prev[2 * q] = __ldg(&k2[m].x);
            prev[2 * q + 1] = __ldg(&k2[m].y);
            ++mine;
        }
    }
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if (lane == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (tid == 0) a.nmatches[p] = total;
}

// -------------------------------------------------------------------------------------------------- SearchByBoW
// src/ORBmatcher.cc:270-404 (key frame against a frame) and :635-768 (key frame against key frame) over flattened
// DBoW2 feature vectors.  A feature sits in exactly one vocabulary node, so the reference's one sequential rule -- a
// matched candidate is skipped by later queries -- never crosses a node: nodes run in parallel, one warp per node of
// frame 1 (binary search for the same id in frame 2: the reference's merge with lower_bound visits exactly the ids
// both maps hold), the node's queries in order inside the warp, lanes over the node's candidates.
// kTri: ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:770-935): same walk over the common nodes, but a candidate
// must also pass the epipole distance test (:866-872) and CheckDistEpipolarLine (:251-268), there is no ratio test, the
// LAST of several equally near candidates wins (`dist > bestDist` skips, :859) and nothing is ever marked matched
// (the reference never sets vbMatched2), so the queries are independent.
template <bool kTri>
__global__ void __launch_bounds__(256)
bow_match_kernel(const BowSearchArgs a) {
    const int lane = threadIdx.x & 31, node = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (node >= a.nn1) return;
    const int id = __ldg(a.node_id1 + node);
    int lo = 0, hi = a.nn2;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(a.node_id2 + mid) < id) lo = mid + 1; else hi = mid;
    }
    if (lo == a.nn2 || __ldg(a.node_id2 + lo) != id) return;
    const int b0 = __ldg(a.node_start2 + lo), b1 = __ldg(a.node_start2 + lo + 1);
    const uint4* d1 = reinterpret_cast<const uint4*>(a.d1);
    const uint4* d2 = reinterpret_cast<const uint4*>(a.d2);
    volatile uint8_t* matched2 = a.matched2;
    constexpr int kListBits = 22;                                       // position inside the node's candidate list
    for (int i = __ldg(a.node_start1 + node); i < __ldg(a.node_start1 + node + 1); ++i) {
        const int idx1 = __ldg(a.node_idx1 + i);
        if (a.valid1 && !__ldg(a.valid1 + idx1)) continue;              // :300-306, :679-684
        const uint4 a0 = __ldg(d1 + 2 * (size_t)idx1), a1 = __ldg(d1 + 2 * (size_t)idx1 + 1);
        uint32_t best = kNone, second = kNone;
        float la = 0.f, lb = 0.f, lc = 0.f;
        bool stereo1 = false;
        if (kTri) {                                                      // epipolar line of kp1 in image 2, :254-256
            const float x1 = __ldg(&a.k1[idx1].x), y1 = __ldg(&a.k1[idx1].y);
            la = __fadd_rn(__fadd_rn(__fmul_rn(x1, a.F12[0]), __fmul_rn(y1, a.F12[3])), a.F12[6]);
            lb = __fadd_rn(__fadd_rn(__fmul_rn(x1, a.F12[1]), __fmul_rn(y1, a.F12[4])), a.F12[7]);
            lc = __fadd_rn(__fadd_rn(__fmul_rn(x1, a.F12[2]), __fmul_rn(y1, a.F12[5])), a.F12[8]);
            stereo1 = a.stereo1 && __ldg(a.stereo1 + idx1);
        }
        for (int b = b0 + lane; b < b1; b += 32) {
            const int idx2 = __ldg(a.node_idx2 + b);
            if ((!kTri && matched2[idx2]) || (a.valid2 && !__ldg(a.valid2 + idx2))) continue;     // :321, :701-706, :844-851
            const int dist = hamming256(a0, a1, __ldg(d2 + 2 * (size_t)idx2), __ldg(d2 + 2 * (size_t)idx2 + 1));
            if (dist >= 256) continue;                                                   // never below the initial 256
            if (kTri) {
                if (dist > a.th_dist) continue;                                          // :859
                const float x2 = __ldg(&a.k2[idx2].x), y2 = __ldg(&a.k2[idx2].y);
                const int o2 = __ldg(&a.k2[idx2].octave) & 31;
                if (!stereo1 && !(a.stereo2 && __ldg(a.stereo2 + idx2))) {               // :866-872
                    const float dx = __fsub_rn(a.ex, x2), dy = __fsub_rn(a.ey, y2);
                    if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.0f, a.sf[o2])) continue;
                }
                const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, x2), __fmul_rn(lb, y2)), lc);      // :258-267
                const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
                if (den == 0.0f) continue;
                const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
                if (!((double)dsqr < __dmul_rn(3.84, (double)a.sig2[o2]))) continue;
                const uint32_t key = ((uint32_t)dist << kListBits) | (uint32_t)(((1 << kListBits) - 1) - (b - b0));   // last of equals
                best = min(best, key);
                continue;
            }
            const uint32_t key = ((uint32_t)dist << kListBits) | (uint32_t)(b - b0);
            second = min(second, max(best, key));
            best = min(best, key);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, best, o), os = __shfl_xor_sync(0xFFFFFFFFu, second, o);
            second = min(min(second, os), max(best, ob));
            best = min(best, ob);
        }
        if (kTri) {
            if (best != kNone && lane == 0)                                              // :878-882
                a.matches12[idx1] = __ldg(a.node_idx2 + b0 + ((1 << kListBits) - 1) - (int)(best & ((1u << kListBits) - 1)));
            continue;
        }
        if (best != kNone && lane == 0) {
            const int bd = (int)(best >> kListBits), sd = second == kNone ? 256 : (int)(second >> kListBits);
            if (bd <= a.th_dist && (float)bd < __fmul_rn(a.ratio, (float)sd)) {          // :337-339, :718-720
                const int idx2 = __ldg(a.node_idx2 + b0 + (int)(best & ((1u << kListBits) - 1)));
                a.matches12[idx1] = idx2;
                matched2[idx2] = 1;
            }
        }
        __syncwarp();
    }
}

// rotation histogram over the matches, three maxima, removal, count (:378-401, :745-766)
__global__ void __launch_bounds__(256)
bow_finish_kernel(const BowSearchArgs a) {
    __shared__ int hist[kHisto];
    __shared__ int keep_bin[3];
    __shared__ int total;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < kHisto) hist[tid] = 0;
    if (tid == 0) total = 0;
    __syncthreads();
    auto rotation_bin = [&](int i1, int i2) {
        float rot = __fsub_rn(__ldg(&a.k1[i1].angle), __ldg(&a.k2[i2].angle));
        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
        const int b = (int)roundf(__fmul_rn(rot, 1.0f / kHisto));
        return b == kHisto ? 0 : b;
    };
    if (a.check_ori) {
        for (int i = tid; i < a.n1; i += 256) {
            const int m = a.matches12[i];
            if (m >= 0) atomicAdd(&hist[rotation_bin(i, m)], 1);
        }
        __syncthreads();
        if (tid == 0) {                                                      // ComputeThreeMaxima, :1723-1764
            int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < kHisto; ++i) {
                const int s = hist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
                else if (s > max3) { max3 = s; i3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { i3 = -1; }
            keep_bin[0] = i1; keep_bin[1] = i2; keep_bin[2] = i3;
        }
        __syncthreads();
    }
    int mine = 0;
    for (int i = tid; i < a.n1; i += 256) {
        int m = a.matches12[i];
        if (m >= 0 && a.check_ori) {
            const int b = rotation_bin(i, m);
            if (b != keep_bin[0] && b != keep_bin[1] && b != keep_bin[2]) { a.matches12[i] = -1; m = -1; }
        }
        mine += m >= 0;
    }
    mine = __reduce_add_sync(0xFFFFFFFFu, mine);
    if (lane == 0 && mine) atomicAdd(&total, mine);
    __syncthreads();
    if (tid == 0) *a.nmatches = total;
}

// --------------------------------------------------------------------------------------------------- launchers
int gridmatch_max_keys() { return 1 << kPosBits; }
int gridmatch_cells() { return kGCells; }


void launch_grid_build(const gforb_keypoint* keys, const int32_t* counts, int n_fixed, int key_stride, int nframes,
                       const float* bounds, int32_t* cell_start, int32_t* cell_idx, cudaStream_t s) {
    const float w_inv = 64.0f / (bounds[1] - bounds[0]), h_inv = 48.0f / (bounds[3] - bounds[2]);   // src/Frame.cc:129-130
    const int ncap = std::max(counts ? key_stride : n_fixed, 2);
    static size_t have[64] = {};
    opt_in_smem(grid_build_kernel, (size_t)4 << kPosBits, have);
    grid_build_kernel<<<nframes, kGridThreads, (size_t)ncap * 4, s>>>(keys, counts, n_fixed, key_stride, bounds[0], bounds[2],
                                                                      w_inv, h_inv, cell_start, cell_idx, ncap);
}

void launch_area_query(const gforb_keypoint* keys, const float* bounds, const int32_t* cell_start, const int32_t* cell_idx,
                       const void* queries, int nq, int32_t* out_count, const int32_t* out_start, int32_t* out_idx,
                       cudaStream_t s) {
    if (nq <= 0) return;
    GridView g{cell_start, cell_idx, bounds[0], bounds[2], 64.0f / (bounds[1] - bounds[0]), 48.0f / (bounds[3] - bounds[2])};
    area_query_kernel<<<(nq + 7) / 8, 256, 0, s>>>(keys, g, static_cast<const AreaQuery*>(queries), nq, out_count, out_start, out_idx);
}

int launch_init_search(const InitSearchArgs& a, const float* bounds, int npairs, cudaStream_t s) {
    InitSearchArgs b = a;
    b.min_x = bounds[0]; b.min_y = bounds[2];
    b.w_inv = 64.0f / (bounds[1] - bounds[0]); b.h_inv = 48.0f / (bounds[3] - bounds[2]);
    const int qcap = std::max(a.n1 ? a.stride1 : a.n1_fixed, 1);
    init_scan_kernel<<<dim3((qcap + 7) / 8, npairs), 256, 0, s>>>(b);
    const int topk = qcap <= 4096 ? kTopK : kTopK / 2;
    const size_t smem = (size_t)qcap * topk * 4 + (4u << kPosBits);
    static size_t have[64] = {};
    if (opt_in_smem(init_resolve_kernel, smem, have)) return -1;
    init_resolve_kernel<<<npairs, kSearchThreads, smem, s>>>(b, topk, qcap);
    return 0;
}

int launch_proj_search(const ProjSearchArgs& a, const float* bounds, int nframes, cudaStream_t s) {
    ProjSearchArgs b = a;
    b.min_x = bounds[0]; b.min_y = bounds[2];
    b.w_inv = 64.0f / (bounds[1] - bounds[0]); b.h_inv = 48.0f / (bounds[3] - bounds[2]);
    const int qcap = a.nq ? a.q_stride : a.nq_fixed;
    if (qcap > 0) proj_scan_kernel<<<dim3((qcap + 7) / 8, nframes), 256, 0, s>>>(b);
    const int topk = qcap <= 4096 ? kTopK : kTopK / 2;           // lists stay within 128 KB of shared memory
    const size_t smem = (size_t)std::max(qcap, 1) * topk * 4 + (4u << kPosBits);
    static size_t have[64] = {};                                 // static shared memory takes 24 KB of the un-opted 48 KB
    if (opt_in_smem(proj_resolve_kernel, smem, have)) return -1;
    proj_resolve_kernel<<<nframes, kSearchThreads, smem, s>>>(b, topk, std::max(qcap, 1));
    return 0;
}

// ORBmatcher::SearchBySim3, the agreement test (src/ORBmatcher.cc:1420-1435): map point i1 of KF1 found key idx2 in KF2
// and map point idx2 of KF2 found key i1 in KF1.  match1 / match2 are the two searches' query_key arrays.
__global__ void __launch_bounds__(256)
sim3_agree_kernel(const int32_t* __restrict__ match1, int n1, const int32_t* __restrict__ match2, int n2,
                  int32_t* __restrict__ match12, int32_t* __restrict__ nfound) {
    int found = 0;
    for (int i1 = blockIdx.x * blockDim.x + threadIdx.x; i1 < n1; i1 += gridDim.x * blockDim.x) {
        const int idx2 = match1[i1];
        const bool ok = idx2 >= 0 && idx2 < n2 && match2[idx2] == i1;
        match12[i1] = ok ? idx2 : -1;
        found += ok;
    }
    found = __reduce_add_sync(0xffffffffu, found);
    if ((threadIdx.x & 31) == 0 && found) atomicAdd(nfound, found);
}

void launch_sim3_agree(const int32_t* match1, int n1, const int32_t* match2, int n2, int32_t* match12, int32_t* nfound,
                       cudaStream_t s) {
    cudaMemsetAsync(nfound, 0, 4, s);
    if (n1 > 0) sim3_agree_kernel<<<std::min((n1 + 255) / 256, 64), 256, 0, s>>>(match1, n1, match2, n2, match12, nfound);
}

void launch_bow_search(const BowSearchArgs& a, cudaStream_t s) {
    if (a.n1 > 0) cudaMemsetAsync(a.matches12, 0xFF, (size_t)a.n1 * 4, s);
    if (a.n2 > 0) cudaMemsetAsync(a.matched2, 0, (size_t)a.n2, s);
    if (a.nn1 > 0 && a.nn2 > 0) {
        if (a.tri) bow_match_kernel<true><<<(a.nn1 + 7) / 8, 256, 0, s>>>(a);
        else bow_match_kernel<false><<<(a.nn1 + 7) / 8, 256, 0, s>>>(a);
    }
    bow_finish_kernel<<<1, 256, 0, s>>>(a);
}

}  // namespace gforb
