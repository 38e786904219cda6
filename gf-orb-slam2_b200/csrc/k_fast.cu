// Per-cell FAST-9/16 with threshold retry: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (reference src/ORBextractor.cc:767-831) over cv::FAST semantics (OpenCV fast.cpp FAST_t<16>,
// fast_score.cpp cornerScore<16>; in-tree restatement src/FAST_NEON.cc:16-66,91-287).
//
// One CTA per (image, cell).  The reference calls cv::FAST(cell, iniTh) and, if that returns NO keypoint
// after non-max suppression, cv::FAST(cell, minTh).  Both calls are folded into one pass:
//   m(p)      = max over the 16 arcs of 9 contiguous circle pixels of min |v - p_k| (same-sign arcs only)
//   corner@t  <=> m > t,  cornerScore = m - 1  (independent of t for a corner)
// so with s = (m > tlo ? m-1 : 0), tlo = min(iniTh, minTh):  score@t = (s >= t ? s : 0), and because NMS is a
// strict '>' against the 8 neighbours (non-corners hold 0, the cell's outer ring holds 0: NMS is cell-local),
//   keep@t(p) = keep@tlo(p) && s(p) >= t.
// One NMS pass therefore serves both thresholds; the block votes on "any keep@ini" and falls back to
// keep@min.  Survivors leave in row-major order (the reference's emission order) via a per-row 64-bit
// bitmap + popcount scan -- no atomics-ordered output.
//
// Data movement: the cell tile (<= 65x65 px) is fetched by ONE TMA request (cp.async.bulk.tensor.3d from the
// padded pyramid {pitch, rows, images}) completing on an mbarrier.  Measured on B200: the innermost TMA
// coordinate must be 16-byte aligned (an unaligned x raises "illegal instruction", tests/cpp/tma_probe.cu), so
// the box starts at the 16-byte boundary below the cell and the cell sits `xoff` (0..15) bytes into the tile.
// Per-cell geometry comes from a host-built 32-byte table entry (one broadcast load) instead of per-CTA
// integer divisions.
//
// Compute (instruction-issue bound; v3 = 4 pixels per thread-instruction in the prefilter):
//  (1) prefilter, 4 horizontally adjacent pixels per 32-bit word: VABSDIFF4 against the words 3 rows up/down and
//      3 bytes left/right (funnel shifts), SWAR "> t" on the byte lanes, ((|up|>t or |down|>t) and (|left|>t or
//      |right|>t)).  This is a superset of cv::FAST's high-speed test (a 9-arc contains one pixel of each
//      opposite pair); a thread walks R rows of one word column and keeps <= 32 flag bits in a register.
//  (2) flagged pixels are compacted into a shared list (warp shuffle scan of the per-thread counts);
//  (3) list entries get the exact arc strength: ring pixel q_k -> one IMAD builds the biased pair
//      {v-q+256, q-v+256} as s16x2, then a 3-input packed min/max network (VIMNMX3) serves both polarities;
//  (4) NMS walks the same list on a zero-initialised score map.
// Algorithmic bytes = level pixels read once + 4 B per survivor.
#include "gforb_internal.cuh"
#include "tma.cuh"

namespace gforb {

namespace {
constexpr int kMaxTP = 80;                // largest tile pitch: align16(59 + 6 + 15)
constexpr int kTileRows = kMaxCellDim;    // 66
constexpr int kThreads = 128;
constexpr int kMaxDomain = 59 * 59;

// Bresenham circle r=3 (src/FAST_NEON.cc:3-7)
__device__ __forceinline__ constexpr int circle_dx(int k) {
    constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    return dx[k];
}
__device__ __forceinline__ constexpr int circle_dy(int k) {
    constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return dy[k];
}

// max over the 16 arcs of min over 9 contiguous of (v - q) and of (q - v); rows[j] = pixel row (j - 3) of the ring
__device__ __forceinline__ int fast_arc_strength(const uint8_t* const (&rows)[7], int v) {
    const uint32_t base = ((uint32_t)(256 - v) << 16) | (uint32_t)(v + 256);
    uint32_t p[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const uint32_t q = rows[circle_dy(k) + 3][circle_dx(k)];
        p[k] = q * 0xFFFFu + base;          // {lo: v - q + 256, hi: q - v + 256}, both in [1, 511]: no borrow
    }
    uint32_t m3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m3[k] = __vimin3_s16x2(p[k], p[(k + 1) & 15], p[(k + 2) & 15]);
    uint32_t m9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m9[k] = __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
    uint32_t b0 = __vimax3_s16x2(m9[0], m9[1], m9[2]), b1 = __vimax3_s16x2(m9[3], m9[4], m9[5]);
    uint32_t b2 = __vimax3_s16x2(m9[6], m9[7], m9[8]), b3 = __vimax3_s16x2(m9[9], m9[10], m9[11]);
    uint32_t b4 = __vimax3_s16x2(m9[12], m9[13], m9[14]);
    b0 = __vimax3_s16x2(b0, b1, b2);
    b3 = __vimax3_s16x2(b3, b4, m9[15]);
    b0 = __vmaxs2(b0, b3);
    return max((int)(b0 & 0xFFFFu), (int)(b0 >> 16)) - 256;
}

// byte lanes of x that exceed t: bit 7 of each lane of the result (other bits are junk).  kadd = (0x7f - t) * 0x01010101, t <= 126
__device__ __forceinline__ uint32_t gt_lanes2(uint32_t a, uint32_t b, uint32_t kadd) {
    return ((a & 0x7f7f7f7fu) + kadd) | ((b & 0x7f7f7f7fu) + kadd) | a | b;
}
}  // namespace

struct FastArgs {
    int img0, cells_total, cell_cap, t_ini, t_min;
    const uint4* cell_tab;
    uint32_t* cell_cnt; uint32_t* cell_pts; int32_t* err;
};

__global__ void __launch_bounds__(kThreads)
fast_cells_kernel(const FastArgs a, const __grid_constant__ TileMaps maps) {
    __shared__ __align__(128) uint8_t tile[kTileRows * kMaxTP];
    __shared__ __align__(16) uint8_t sc[kTileRows * kMaxTP + 32];   // corner scores, same coordinates as `tile`
    __shared__ uint16_t list[kMaxDomain + 3];          // (row << 8) | tile column of pixels passing the prefilter
    __shared__ unsigned long long keep_lo[64], keep_hi[64];
    __shared__ int row_base[64];
    __shared__ __align__(8) uint64_t bar;
    __shared__ int s_any_hi, s_nlist;

    const int img = a.img0 + blockIdx.y;
    const size_t out_cell = (size_t)img * a.cells_total + blockIdx.x;
    const uint4 ca = __ldg(a.cell_tab + 2 * blockIdx.x), cb = __ldg(a.cell_tab + 2 * blockIdx.x + 1);
    const int cw = (ca.y >> 8) & 0xFF, ch = (ca.y >> 16) & 0xFF;
    if (cw == 0) {                                      // cell skipped by the reference (:796,805) or degenerate
        if (threadIdx.x == 0) a.cell_cnt[out_cell] = 0;
        return;
    }
    const int xoff = ca.y & 0xFF, l = ca.y >> 24;
    const int tp = ca.w & 0xFF, th = (ca.w >> 8) & 0xFF, R = (ca.w >> 16) & 0xFF, nw = ca.w >> 24;
    const int dw = cw - 6, dh = ch - 6;                 // FAST domain: rows/cols 3 .. dim-4 of the cell image
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        s_any_hi = 0; s_nlist = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, (uint32_t)(tp * th));
        tma_load_tile_3d(tile, &maps.fast[l], (int)(ca.x & 0xFFFF), (int)(ca.x >> 16), img, &bar);
    }
    // while the tile is in flight: clear the score rows NMS will touch and the row bitmaps
    {
        const int n16 = (ch * tp + 31) >> 4;
        for (int i = threadIdx.x; i < n16; i += kThreads) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
        if (threadIdx.x < 64) { keep_lo[threadIdx.x] = 0; keep_hi[threadIdx.x] = 0; }
    }
    const int t_ini = min(max(a.t_ini, 0), 255), t_min = min(max(a.t_min, 0), 255);
    const int tlo = min(t_ini, t_min);
    const uint32_t kadd = (uint32_t)(0x7f - min(tlo, 126)) * 0x01010101u;   // a lower prefilter threshold stays a superset
    const int tpw = tp >> 2;
    const int xlo = xoff + 3, xhi = xlo + dw;           // domain columns [xlo, xhi) in tile bytes
    const int w0 = xlo >> 2;
    const int items = (int)cb.y, rcp = (int)cb.x;       // items = strips x nw; rcp = ceil(65536 / nw)

    mbar_wait(&bar, 0);

    // ---- pass 1: 4-pixel prefilter + compaction ----
    const uint32_t* tw = reinterpret_cast<const uint32_t*>(tile);
    for (int i0 = 0; i0 < items; i0 += kThreads) {
        const int i = i0 + threadIdx.x;
        uint32_t F = 0;                                 // bit 8k + (8 - R) + r  <->  pixel k of the word, strip row r
        int s = 0, w = 0;
        if (i < items) {
            s = (i * rcp) >> 16; w = i - s * nw;
            const int row0 = 3 + s * R;
            const uint32_t* p = tw + row0 * tpw + w0 + w;
            for (int r = 0; r < R; ++r, p += tpw) {
                uint32_t f = 0;
                if (row0 + r < 3 + dh) {
                    const uint32_t C = p[0], Wm = p[-1], Wp = p[1], U = p[-3 * tpw], D = p[3 * tpw];
                    const uint32_t Lw = __funnelshift_r(Wm, C, 8), Rw = __funnelshift_r(C, Wp, 24);
                    const uint32_t v = gt_lanes2(__vabsdiffu4(C, U), __vabsdiffu4(C, D), kadd);
                    const uint32_t h = gt_lanes2(__vabsdiffu4(C, Lw), __vabsdiffu4(C, Rw), kadd);
                    f = v & h & 0x80808080u;
                }
                F = (F >> 1) | f;
            }
            // columns outside [xlo, xhi) (first / last word of the row)
            const int x0 = (w0 + w) << 2;
            uint32_t cm = 0xFFFFFFFFu;
            if (x0 < xlo) cm <<= 8 * (xlo - x0);
            if (x0 + 4 > xhi) cm &= 0xFFFFFFFFu >> (8 * (x0 + 4 - xhi));
            F &= cm;
        }
        const int c = __popc(F);
        int inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        const int tot = __shfl_sync(0xffffffffu, inc, 31);
        if (tot) {
            int base = 0;
            if (lane == 31) base = atomicAdd(&s_nlist, tot);
            base = __shfl_sync(0xffffffffu, base, 31);
            uint16_t* out = list + base + inc - c;
            // entry = (row << 8) | column with row = 3 + s*R + r, r = (bit & 7) - (8 - R), column = x0 + (bit >> 3)
            const int ebase = ((3 + s * R - (8 - R)) << 8) | ((w0 + w) << 2);
            while (F) {
                const int bit = 31 - __clz(F);
                F ^= 1u << bit;
                *out++ = (uint16_t)(ebase + ((bit & 7) << 8) + (bit >> 3));
            }
        }
    }
    __syncthreads();
    const int nlist = s_nlist;

    // ---- pass 2: exact corner strength of the listed pixels ----
    for (int k = threadIdx.x; k < nlist; k += kThreads) {
        const int e = list[k];
        const int off = (e >> 8) * tp + (e & 0xFF);
        const uint8_t* c = tile + off;
        const uint8_t* const rows[7] = {c - 3 * tp, c - 2 * tp, c - tp, c, c + tp, c + 2 * tp, c + 3 * tp};
        const int m = fast_arc_strength(rows, c[0]);
        if (m > tlo) sc[off] = (uint8_t)(m - 1);
    }
    __syncthreads();

    // ---- pass 3: strict 3x3 NMS on the score map (zero outside the domain), again only over the list ----
    for (int k = threadIdx.x; k < nlist; k += kThreads) {
        const int e = list[k];
        const int py = e >> 8, px = e & 0xFF;
        const uint8_t* q = sc + py * tp + px;
        const int s = q[0];
        if (s == 0) continue;
        const bool keep = s > q[-1] && s > q[1] && s > q[-tp - 1] && s > q[-tp] && s > q[-tp + 1] &&
                          s > q[tp - 1] && s > q[tp] && s > q[tp + 1];
        if (!keep) continue;
        const unsigned long long bit = 1ull << (px - xlo);
        atomicOr(&keep_lo[py - 3], bit);
        if (s >= t_ini) { atomicOr(&keep_hi[py - 3], bit); s_any_hi = 1; }
    }
    __syncthreads();

    // keep@ini if it found anything, else keep@min (s >= t_min filter applies when t_min > tlo)
    const bool use_hi = s_any_hi != 0;
    if (threadIdx.x < 64) {
        unsigned long long m = 0;
        if ((int)threadIdx.x < dh) {
            m = use_hi ? keep_hi[threadIdx.x] : keep_lo[threadIdx.x];
            if (!use_hi && t_min > tlo) {   // iniTh < minTh configuration: retry threshold is the higher one
                unsigned long long f = 0;
                for (unsigned long long r = m; r; r &= r - 1) {
                    const int bx = __ffsll((long long)r) - 1;
                    if (sc[(threadIdx.x + 3) * tp + bx + xlo] >= t_min) f |= 1ull << bx;
                }
                m = f;
            }
        }
        keep_lo[threadIdx.x] = m;
        row_base[threadIdx.x] = __popcll(m);
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        // exclusive scan of the <= 59 row counts by one warp
        const int c0 = row_base[lane], c1 = row_base[lane + 32];
        int i0 = c0, i1 = c1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t0 = __shfl_up_sync(0xffffffffu, i0, o), t1 = __shfl_up_sync(0xffffffffu, i1, o);
            if (lane >= o) { i0 += t0; i1 += t1; }
        }
        const int tot0 = __shfl_sync(0xffffffffu, i0, 31), tot1 = __shfl_sync(0xffffffffu, i1, 31);
        row_base[lane] = i0 - c0;
        row_base[lane + 32] = tot0 + i1 - c1;
        if (lane == 0) {
            int total = tot0 + tot1;
            if (total > a.cell_cap) { atomicOr(a.err, kErrCellOverflow); total = 0; }
            a.cell_cnt[out_cell] = (uint32_t)total;
            s_nlist = total;
        }
    }
    __syncthreads();
    if (s_nlist == 0) return;
    if ((int)threadIdx.x < dh) {
        uint32_t* out = a.cell_pts + out_cell * (size_t)a.cell_cap + row_base[threadIdx.x];
        const int yy = threadIdx.x + 3;
        const int ox = (int)(ca.z & 0xFFFF), oy = (int)(ca.z >> 16);      // j*wCell, i*hCell (:824-825)
        for (unsigned long long r = keep_lo[threadIdx.x]; r; r &= r - 1) {
            const int bx = __ffsll((long long)r) - 1;
            *out++ = pack_cand(ox + bx + 3, oy + yy, sc[yy * tp + bx + xlo]);
        }
    }
}

// Host: per-cell table (src/ORBextractor.cc:785-808 geometry), 2 x uint4 per cell in the kernel's cell order.
void build_fast_cell_table(const Geom& g, std::vector<uint32_t>& tab) {
    tab.assign((size_t)g.cells_total * 8, 0u);
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        for (int c = 0; c < L.ncells; ++c) {
            uint32_t* e = &tab[(size_t)(L.cell_base + c) * 8];
            const int ci = c / L.nCols, cj = c - ci * L.nCols;
            const int iniX = kMinBorder + cj * L.wCell, iniY = kMinBorder + ci * L.hCell;
            // src/ORBextractor.cc:796,805: asymmetric skip tests, as written
            const int cw = std::min(iniX + L.wCell + 6, L.maxBX) - iniX;
            const int ch = std::min(iniY + L.hCell + 6, L.maxBY) - iniY;
            const int dw = cw - 6, dh = ch - 6;
            if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6 || dw <= 0 || dh <= 0) continue;   // cw = 0: skipped
            const int gx = kXPad + iniX, xoff = gx & 15;
            const int xlo = xoff + 3, xhi = xlo + dw;
            const int nw = ((xhi - 1) >> 2) - (xlo >> 2) + 1;
            const int R = std::min(std::max((dh * nw + kThreads - 1) / kThreads, 1), 8);
            const int strips = (dh + R - 1) / R;
            e[0] = (uint32_t)(gx - xoff) | ((uint32_t)(kEdge + iniY) << 16);
            e[1] = (uint32_t)xoff | ((uint32_t)cw << 8) | ((uint32_t)ch << 16) | ((uint32_t)l << 24);
            e[2] = (uint32_t)(cj * L.wCell) | ((uint32_t)(ci * L.hCell) << 16);
            e[3] = (uint32_t)L.fast_bw | ((uint32_t)L.fast_bh << 8) | ((uint32_t)R << 16) | ((uint32_t)nw << 24);
            e[4] = (uint32_t)((65536 + nw - 1) / nw);
            e[5] = (uint32_t)(strips * nw);
        }
    }
}

void launch_fast(const Geom& g, const TileMaps& m, const Buffers& b, int nimg, cudaStream_t s) {
    FastArgs a;
    a.img0 = g.img0; a.cells_total = g.cells_total; a.cell_cap = g.cell_cap; a.t_ini = g.ini_th; a.t_min = g.min_th;
    a.cell_tab = reinterpret_cast<const uint4*>(b.cell_tab);
    a.cell_cnt = b.cell_cnt; a.cell_pts = b.cell_pts; a.err = b.err;
    dim3 grd(g.cells_total, nimg);
    fast_cells_kernel<<<grd, kThreads, 0, s>>>(a, m);
}

}  // namespace gforb
