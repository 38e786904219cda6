// Quadtree keypoint culling: ORBextractor::DistributeOctTree + ExtractorNode::DivideNode
// (reference src/ORBextractor.cc:481-763), reformulated for one CTA per (image, level).
//
// The reference walks a std::list, splitting nodes one at a time and copying KeyPoint vectors.  What its
// result depends on is only (a) which node every point ends up in, (b) the order of the node list, and
// (c) the order in which nodes are split once the "largest first" regime starts.  Reformulation:
//   * points never move: each carries a node index; original candidate order is the index order, so
//     "first strict max response" per node (:744-760) is an atomicMax on response + atomicMin on index;
//   * the list is a table kept IN LIST ORDER.  A pass that creates children c_0..c_{C-1} (in creation
//     order) and leaves U nodes unsplit yields the list [c_{C-1} .. c_0, unsplit nodes in old order],
//     because children are push_front'ed and parents erased in place (:606-665, :688-727);
//   * a pass = parallel speculative split of every splittable node (per-quadrant counts by smem atomics),
//     then scans decide creation indices.  In the regular regime every node with >1 point splits, in list
//     order.  In the "largest first" regime (:673-737) candidates are ordered by (count desc, newest
//     first) and only the prefix up to the split that reaches N is committed.  The reference breaks count
//     ties by heap address (:591,684); frozen as "newest first" = smaller table index.
// Latency-bound bookkeeping: reported by time, not by a bandwidth roofline.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kT = 256;

struct Smem {
    uint32_t* pts;      // [max_cand] packed x|y|resp
    uint16_t* pnode;    // [max_cand]
    short4* box[2];     // [NC] x0,x1,y0,y1 (ping-pong)
    int* cnt[2];        // [NC]
    int* cnt4;          // [4*NC] child counts, later best response
    int* remap;         // [4*NC] child table index, later best index
    int* keepmap;       // [NC]
    int* a;             // [NC] scan scratch 1 (children per node / creation prefix)
    int* b;             // [NC] scan scratch 2 (unsplit prefix / sorted increments)
    uint32_t* key;      // [NC] sort keys
    int* order;         // [NC] node index by sorted rank
    uint8_t* split;     // [NC]
    int* warp;          // [128] scan scratch
};

__device__ __forceinline__ Smem carve(uint8_t* base, int max_cand, int NC) {
    Smem s;
    size_t o = 0;
    auto take = [&](size_t bytes) { uint8_t* p = base + o; o += (bytes + 15) & ~(size_t)15; return p; };
    s.pts = (uint32_t*)take(4 * (size_t)max_cand);
    s.pnode = (uint16_t*)take(2 * (size_t)max_cand);
    s.box[0] = (short4*)take(8 * (size_t)NC); s.box[1] = (short4*)take(8 * (size_t)NC);
    s.cnt[0] = (int*)take(4 * (size_t)NC);    s.cnt[1] = (int*)take(4 * (size_t)NC);
    s.cnt4 = (int*)take(16 * (size_t)NC);
    s.remap = (int*)take(16 * (size_t)NC);
    s.keepmap = (int*)take(4 * (size_t)NC);
    s.a = (int*)take(4 * (size_t)NC);
    s.b = (int*)take(4 * (size_t)NC);
    s.key = (uint32_t*)take(4 * (size_t)NC);
    s.order = (int*)take(4 * (size_t)NC);
    s.split = (uint8_t*)take((size_t)NC);
    s.warp = (int*)take(4 * 128);      // [0, 64): block_scan_pair's two buffers; [64, 104): block_excl_scan
    // the pointers come out of a struct: without the hint every access through them is a generic LD / ST / ATOM
    __builtin_assume(__isShared(s.pts));   __builtin_assume(__isShared(s.pnode));  __builtin_assume(__isShared(s.box[0]));
    __builtin_assume(__isShared(s.box[1])); __builtin_assume(__isShared(s.cnt[0])); __builtin_assume(__isShared(s.cnt[1]));
    __builtin_assume(__isShared(s.cnt4));  __builtin_assume(__isShared(s.remap));  __builtin_assume(__isShared(s.keepmap));
    __builtin_assume(__isShared(s.a));     __builtin_assume(__isShared(s.b));      __builtin_assume(__isShared(s.key));
    __builtin_assume(__isShared(s.order)); __builtin_assume(__isShared(s.split));  __builtin_assume(__isShared(s.warp));
    return s;
}

// in-place exclusive scan of v[0..n) by the whole block; returns the total.  scratch: >= 34 ints.
__device__ int block_excl_scan(int* v, int n, int* scratch) {
    const int per = (n + kT - 1) / kT;
    const int lo = min((int)threadIdx.x * per, n), hi = min(lo + per, n);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += v[i];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                       // protects scratch reuse across calls
    if (lane == 31) scratch[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < kT / 32 ? scratch[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < kT / 32) scratch[lane] = winc - w;
        if (lane == kT / 32 - 1) scratch[33] = winc;
    }
    __syncthreads();
    int run = scratch[wid] + inc - sum;
    for (int i = lo; i < hi; ++i) { const int t = v[i]; v[i] = run; run += t; }
    const int total = scratch[33];
    __syncthreads();
    return total;
}

// One value per thread: exclusive prefix of v over the block, the block totals of v and of a second value w, ONE barrier.
// scratch holds two buffers of 2 * kT / 32 ints used alternately (parity), so a warp that runs ahead into the next call
// writes the other buffer while slower warps still read this one; a third call cannot start before every warp has
// passed the second call's barrier.
__device__ __forceinline__ int block_scan_pair(int v, int w, int* scratch, int& parity, int& v_total, int& w_total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    constexpr int kW = kT / 32;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    const int wsum = __reduce_add_sync(0xffffffffu, w);
    int* sc = scratch + parity * 2 * kW;
    parity ^= 1;
    if (lane == 31) { sc[wid] = inc; sc[kW + wid] = wsum; }
    __syncthreads();
    const int x = lane < kW ? sc[lane] : 0, y = lane < kW ? sc[kW + lane] : 0;
    int xinc = x;
#pragma unroll
    for (int o = 1; o < kW; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, xinc, o);
        if (lane >= o) xinc += t;
    }
    v_total = __shfl_sync(0xffffffffu, xinc, kW - 1);
    w_total = __reduce_add_sync(0xffffffffu, y);
    return __shfl_sync(0xffffffffu, xinc - x, wid) + inc - v;
}

__device__ __forceinline__ int quadrant(uint32_t p, short4 bx) {
    const int sx = bx.x + ((bx.y - bx.x + 1) >> 1);   // UL.x + ceil((UR.x-UL.x)/2)   (:483)
    const int sy = bx.z + ((bx.w - bx.z + 1) >> 1);   // UL.y + ceil((BR.y-UL.y)/2)   (:484)
    return (cand_x(p) >= sx ? 1 : 0) + (cand_y(p) >= sy ? 2 : 0);   // n1,n2,n3,n4 (:514-527)
}
__device__ __forceinline__ short4 child_box(short4 bx, int q) {
    const short sx = (short)(bx.x + ((bx.y - bx.x + 1) >> 1));
    const short sy = (short)(bx.z + ((bx.w - bx.z + 1) >> 1));
    short4 c;
    c.x = (q & 1) ? sx : bx.x;  c.y = (q & 1) ? bx.y : sx;
    c.z = (q & 2) ? sy : bx.z;  c.w = (q & 2) ? bx.w : sy;
    return c;
}
}  // namespace

namespace {
size_t smem_bytes_for(int cap, int max_nodes) {
    auto r = [](size_t b) { return (b + 15) & ~(size_t)15; };
    const size_t mc = (size_t)cap, nc = (size_t)max_nodes;
    return r(4 * mc) + r(2 * mc) + 2 * r(8 * nc) + 2 * r(4 * nc) + 2 * r(16 * nc) + 5 * r(4 * nc) + r(nc) + r(512);
}
}  // namespace

// Two tiers: every (image, level) first runs with room for kSmallCap candidates (about 44 KB of shared memory: five
// CTAs per SM instead of two -- the kernel is latency-bound, so residency is throughput); the rare pair that holds
// more flags itself and is redone by a second launch with the full max_cand working set.
#ifndef GFORB_OCT_SMALLCAP
#define GFORB_OCT_SMALLCAP 4096
#endif
#ifndef GFORB_OCT_MINB
#define GFORB_OCT_MINB 5
#endif
constexpr int kSmallCap = GFORB_OCT_SMALLCAP;
int octree_small_cap(const Geom& g) { return g.max_cand > kSmallCap ? kSmallCap : g.max_cand; }
size_t octree_smem_bytes(const Geom& g) { return smem_bytes_for(g.max_cand, g.max_nodes); }

// one (image, level) pair by the whole CTA; every return is block-uniform
__device__ __forceinline__ void octree_pair(const Geom& g, const uint32_t* __restrict__ cell_cnt,
                                            const uint32_t* __restrict__ cell_pts, int32_t* __restrict__ lvl_cnt,
                                            uint32_t* __restrict__ lvl_pts, int32_t* __restrict__ err,
                                            int32_t* __restrict__ retry, int cap, int l, int img) {
    extern __shared__ __align__(16) uint8_t smem_raw[];      // declared here, not passed in: keeps every access an LDS / ATOMS
    int32_t* my_retry = retry + (size_t)img * g.nlevels + l;
    const LevelGeom& L = g.lv[l];
    const int NC = g.max_nodes;
    Smem s = carve(smem_raw, cap, NC);
    __shared__ int sh_n, sh_flag;

    // ---- gather this level's FAST survivors in reference order (cell row-major, row-major inside) ----
    const uint32_t* ccnt = cell_cnt + (size_t)img * g.cells_total + L.cell_base;
    const uint32_t* cpts = cell_pts + ((size_t)img * g.cells_total + L.cell_base) * (size_t)g.cell_cap;
    int n = 0, parity = 0;
    for (int c0 = 0; c0 < L.ncells; c0 += kT) {
        const int c = c0 + threadIdx.x;
        const int mine = c < L.ncells ? (int)ccnt[c] : 0;
        int tot, unused;
        const int at = n + block_scan_pair(mine, 0, s.warp, parity, tot, unused);      // one barrier per 256 cells
        if (n + tot <= cap)
            for (int k = 0; k < mine; ++k) s.pts[at + k] = cpts[(size_t)c * g.cell_cap + k];
        n += tot;
    }
    __syncthreads();
    int32_t* out_cnt = lvl_cnt + (size_t)img * g.nlevels + l;
    uint32_t* out_pts = lvl_pts + (size_t)img * g.kstage + L.kbase;
    if (n > cap) {
        if (threadIdx.x == 0) {
            if (cap < g.max_cand) *my_retry = 1;          // redo with the full working set (tier 1)
            else { atomicOr(err, kErrCandOverflow); *out_cnt = 0; }
        }
        return;
    }
    if (n == 0 || L.nIni <= 0) {
        if (threadIdx.x == 0) *out_cnt = 0;
        return;
    }

    // ---- initial nodes (:543-596) ----
    const int N = L.quota;
    const int H = L.maxBY - kMinBorder;           // maxY - minY
    int cur = 0, S = 0;
    for (int i = threadIdx.x; i < L.nIni; i += kT) {
        short4 bx;
        bx.x = (short)(int)__fmul_rn(L.hX, (float)i);
        bx.y = (short)(int)__fmul_rn(L.hX, (float)(i + 1));
        bx.z = 0; bx.w = (short)H;
        s.box[1][i] = bx;          // staged in the other table, compacted below
        s.cnt[1][i] = 0;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += kT) {
        const int slot = (int)__fdiv_rn((float)cand_x(s.pts[p]), L.hX);   // vpIniNodes[kp.pt.x/hX] (:569)
        s.pnode[p] = (uint16_t)slot;
        atomicAdd(&s.cnt[1][slot], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < L.nIni; i += kT) s.a[i] = s.cnt[1][i] > 0 ? 1 : 0;
    __syncthreads();
    S = block_excl_scan(s.a, L.nIni, s.warp + 64);// empty initial nodes are erased (:585-586)
    for (int i = threadIdx.x; i < L.nIni; i += kT)
        if (s.cnt[1][i] > 0) { s.box[0][s.a[i]] = s.box[1][i]; s.cnt[0][s.a[i]] = s.cnt[1][i]; }
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += kT) s.pnode[p] = (uint16_t)s.a[s.pnode[p]];
    __syncthreads();

    // ---- split passes ----
    // Barriers are what this kernel waits on (ncu: 7 barrier-stall cycles per issued instruction), so a pass of the regular
    // regime is four of them: count, one scan that carries (children before | unsplit before | nodes to expand) in a
    // single sweep, build, move.  The child counters are cleared by the move step of the pass before.
    bool largest_first = false;
    for (int i = threadIdx.x; i < 4 * S; i += kT) s.cnt4[i] = 0;
    __syncthreads();
    for (;;) {
        // selects, not s.box[cur]: a dynamically indexed member sends the whole pointer struct to local memory, and every access
        // through a pointer re-loaded from there is a generic LD / ST / ATOM instead of LDS / STS / ATOMS
        short4* box = cur ? s.box[1] : s.box[0]; int* cnt = cur ? s.cnt[1] : s.cnt[0];
        short4* nbox = cur ? s.box[0] : s.box[1]; int* ncnt = cur ? s.cnt[0] : s.cnt[1];

        // A. speculative child counts of every splittable node
        for (int p = threadIdx.x; p < n; p += kT) {
            const int nd = s.pnode[p];
            if (cnt[nd] > 1) atomicAdd(&s.cnt4[4 * nd + quadrant(s.pts[p], box[nd])], 1);
        }
        __syncthreads();

        int C, S_new, nToExpand;
        if (!largest_first) {
            // B + C. every node with more than one point splits, in list order (:606-665): a thread owns a contiguous run
            // of nodes, the scan gives the children created / the unsplit nodes before its run
            const int per = (S + kT - 1) / kT;
            const int i0 = min((int)threadIdx.x * per, S), i1 = min(i0 + per, S);
            int nch = 0, uns = 0, nexp = 0;
            for (int i = i0; i < i1; ++i) {
                if (cnt[i] > 1) {
                    const int* c4 = s.cnt4 + 4 * i;
                    nch += (c4[0] > 0) + (c4[1] > 0) + (c4[2] > 0) + (c4[3] > 0);
                    nexp += (c4[0] > 1) + (c4[1] > 1) + (c4[2] > 1) + (c4[3] > 1);
                } else ++uns;
            }
            int tot;
            const int pre = block_scan_pair(nch | (uns << 16), nexp, s.warp, parity, tot, nToExpand);
            C = tot & 0xFFFF;
            S_new = C + (tot >> 16);
            if (S_new > NC) {
                if (threadIdx.x == 0) { atomicOr(err, kErrNodeOverflow); *out_cnt = 0; }
                return;
            }
            int c = pre & 0xFFFF, u = pre >> 16;
            for (int i = i0; i < i1; ++i) {
                const bool sp = cnt[i] > 1;
                s.split[i] = sp;
                if (sp) {
                    const short4 bx = box[i];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int k = s.cnt4[4 * i + q];
                        if (k > 0) {
                            const int idx = C - 1 - c;
                            nbox[idx] = child_box(bx, q);
                            ncnt[idx] = k;
                            s.remap[4 * i + q] = idx;
                            ++c;
                        }
                    }
                } else {
                    const int idx = C + u;
                    nbox[idx] = box[i];
                    ncnt[idx] = cnt[i];
                    s.keepmap[i] = idx;
                    ++u;
                }
            }
            __syncthreads();
        } else {
            // candidates = nodes with >1 point (all were created by the previous pass), ordered by
            // (count desc, newest = smaller index first)  (:684-686)
            int nsplit;
            for (int i = threadIdx.x; i < S; i += kT) {
                s.key[i] = cnt[i] > 1 ? ((uint32_t)cnt[i] << 16) | (uint32_t)(0xFFFF - i) : 0u;
                s.split[i] = 0;
            }
            __syncthreads();
            for (int i = threadIdx.x; i < S; i += kT) {
                const uint32_t k = s.key[i];
                if (k == 0) continue;
                int rank = 0;
                for (int j = 0; j < S; ++j) rank += s.key[j] > k;
                s.order[rank] = i;
            }
            if (threadIdx.x == 0) sh_n = 0;
            __syncthreads();
            int m_local = 0;
            for (int i = threadIdx.x; i < S; i += kT) m_local += s.key[i] != 0;
            if (m_local) atomicAdd(&sh_n, m_local);
            __syncthreads();
            const int M = sh_n;
            // size after committing sorted candidates 0..r = S + sum_{k<=r} (children_k - 1)
            for (int r = threadIdx.x; r < M; r += kT) {
                const int* c4 = s.cnt4 + 4 * s.order[r];
                const int nch = (c4[0] > 0) + (c4[1] > 0) + (c4[2] > 0) + (c4[3] > 0);
                s.a[r] = nch;
                s.b[r] = nch - 1;
            }
            if (threadIdx.x == 0) sh_flag = M;            // first rank whose commit reaches N
            __syncthreads();
            block_excl_scan(s.b, M, s.warp + 64);
            for (int r = threadIdx.x; r < M; r += kT) {
                const int after = S + s.b[r] + (s.a[r] - 1);
                if (after >= N) atomicMin(&sh_flag, r);
            }
            __syncthreads();
            nsplit = min(sh_flag + 1, M);                 // break right after the split that reaches N (:729-730)
            for (int r = threadIdx.x; r < M; r += kT) if (r >= nsplit) s.a[r] = 0;
            __syncthreads();
            C = block_excl_scan(s.a, M, s.warp + 64);     // a[r] = children created before rank r's
            // scatter creation prefix back to node indexing (keepmap doubles as scratch here)
            for (int r = threadIdx.x; r < nsplit; r += kT) { s.split[s.order[r]] = 1; s.keepmap[s.order[r]] = s.a[r]; }
            __syncthreads();
            for (int i = threadIdx.x; i < S; i += kT) { s.a[i] = s.split[i] ? s.keepmap[i] : 0; s.b[i] = s.split[i] ? 0 : 1; }
            __syncthreads();
            const int U = block_excl_scan(s.b, S, s.warp + 64);    // b[i] = unsplit nodes before i
            S_new = C + U;
            if (S_new > NC) {
                if (threadIdx.x == 0) { atomicOr(err, kErrNodeOverflow); *out_cnt = 0; }
                return;
            }
            // C. build the next table in list order: children reversed, then unsplit nodes in old order
            if (threadIdx.x == 0) sh_n = 0;
            __syncthreads();
            int expand_local = 0;
            for (int i = threadIdx.x; i < S; i += kT) {
                if (s.split[i]) {
                    int c = s.a[i];
                    const short4 bx = box[i];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int k = s.cnt4[4 * i + q];
                        if (k > 0) {
                            const int idx = C - 1 - c;
                            nbox[idx] = child_box(bx, q);
                            ncnt[idx] = k;
                            s.remap[4 * i + q] = idx;
                            expand_local += k > 1;
                            ++c;
                        }
                    }
                } else {
                    const int idx = C + s.b[i];
                    nbox[idx] = box[i];
                    ncnt[idx] = cnt[i];
                    s.keepmap[i] = idx;
                }
            }
            if (expand_local) atomicAdd(&sh_n, expand_local);
            __syncthreads();
            nToExpand = sh_n;
        }
        // D. move the points; clear the child counters of the next pass (the old ones have been consumed)
        for (int p = threadIdx.x; p < n; p += kT) {
            const int nd = s.pnode[p];
            s.pnode[p] = (uint16_t)(s.split[nd] ? s.remap[4 * nd + quadrant(s.pts[p], box[nd])] : s.keepmap[nd]);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 4 * S_new; i += kT) s.cnt4[i] = 0;
        // (no barrier: the next reader of cnt4 is pass A's atomicAdd, behind the barrier below or the loop's first one)

        // E. termination (:667-672, :733-735)
        const int prevSize = S;
        S = S_new;
        cur ^= 1;
        if (S >= N || S == prevSize) break;
        if (!largest_first && S + 3 * nToExpand > N) largest_first = true;
        __syncthreads();
    }
    __syncthreads();

    // ---- best response per node, first in candidate order wins ties (:744-760) ----
    int* best_r = s.cnt4; int* best_i = s.remap;
    for (int i = threadIdx.x; i < S; i += kT) { best_r[i] = -1; best_i[i] = 0x7fffffff; }
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += kT) atomicMax(&best_r[s.pnode[p]], cand_r(s.pts[p]));
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += kT)
        if (cand_r(s.pts[p]) == best_r[s.pnode[p]]) atomicMin(&best_i[s.pnode[p]], p);
    __syncthreads();
    if (S > L.kcap) {
        if (threadIdx.x == 0) { atomicOr(err, kErrKpOverflow); *out_cnt = 0; }
        return;
    }
    for (int i = threadIdx.x; i < S; i += kT) out_pts[i] = s.pts[best_i[i]];
    if (threadIdx.x == 0) *out_cnt = S;
}

__global__ void __launch_bounds__(kT, GFORB_OCT_MINB)
octree_kernel(const Geom g, const uint32_t* __restrict__ cell_cnt, const uint32_t* __restrict__ cell_pts,
              int32_t* __restrict__ lvl_cnt, uint32_t* __restrict__ lvl_pts, int32_t* __restrict__ err,
              int32_t* __restrict__ retry, int cap) {
    // grid = (images, levels): the block scheduler hands out x first, so the large levels of every image start before the
    // small ones (longest job first: the launch's tail is a level-7 CTA, not a level-0 one)
    octree_pair(g, cell_cnt, cell_pts, lvl_cnt, lvl_pts, err, retry, cap, blockIdx.y, g.img0 + blockIdx.x);
}

// Second tier: the pairs the first one could not hold, with the full working set.  Almost always there are none, so the
// launch is a few hundred CTAs that each look at a slice of the retry flags in parallel (one flag per thread) and leave;
// one CTA per pair with the full shared-memory footprint cost 12 us per step just to read 4,096 flags two CTAs per SM at a time.
__global__ void __launch_bounds__(kT, 1)
octree_redo_kernel(const Geom g, const uint32_t* __restrict__ cell_cnt, const uint32_t* __restrict__ cell_pts,
                   int32_t* __restrict__ lvl_cnt, uint32_t* __restrict__ lvl_pts, int32_t* __restrict__ err,
                   int32_t* __restrict__ retry, int cap, int npairs, int per) {
    __shared__ int s_todo[kT], s_ntodo;
    if (threadIdx.x == 0) s_ntodo = 0;
    __syncthreads();
    const int p = blockIdx.x * per + threadIdx.x;            // per <= kT
    if ((int)threadIdx.x < per && p < npairs) {
        int32_t* f = retry + (size_t)g.img0 * g.nlevels + p;
        if (*f) { *f = 0; s_todo[atomicAdd(&s_ntodo, 1)] = p; }
    }
    __syncthreads();
    const int ntodo = s_ntodo;
    for (int i = 0; i < ntodo; ++i) {
        const int q = s_todo[i];
        octree_pair(g, cell_cnt, cell_pts, lvl_cnt, lvl_pts, err, retry, cap, q % g.nlevels, g.img0 + q / g.nlevels);
        __syncthreads();                                     // the working set is reused by the next pair
    }
}

int octree_configure(const Geom& g) {
    static size_t have[64] = {};       // high-water mark: engines of different footprints share the kernel symbol
    static size_t have_redo[64] = {};
    if (opt_in_smem(octree_kernel, smem_bytes_for(octree_small_cap(g), g.max_nodes), have)) return (int)cudaErrorInvalidValue;
    return opt_in_smem(octree_redo_kernel, octree_smem_bytes(g), have_redo) ? (int)cudaErrorInvalidValue : 0;
}

void launch_octree(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd(nimg, g.nlevels);
    const int small = octree_small_cap(g);
    octree_kernel<<<grd, kT, smem_bytes_for(small, g.max_nodes), s>>>(g, b.cell_cnt, b.cell_pts, b.lvl_cnt, b.lvl_pts, b.err,
                                                                    b.oct_retry, small);
    if (small < g.max_cand) {
        const int npairs = nimg * g.nlevels;
        const int ctas = std::max(std::min(npairs, 296), (npairs + kT - 1) / kT);      // about two per SM; a slice <= kT flags
        const int per = (npairs + ctas - 1) / ctas;
        octree_redo_kernel<<<ctas, kT, octree_smem_bytes(g), s>>>(g, b.cell_cnt, b.cell_pts, b.lvl_cnt, b.lvl_pts, b.err,
                                                                 b.oct_retry, g.max_cand, npairs, per);
    }
}

}  // namespace gforb
