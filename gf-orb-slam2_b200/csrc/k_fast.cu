// Per-cell FAST-9/16 with threshold retry: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (reference src/ORBextractor.cc:767-831) over cv::FAST semantics (OpenCV fast.cpp FAST_t<16>,
// fast_score.cpp cornerScore<16>; in-tree restatement src/FAST_NEON.cc:16-66,91-287).
//
// One WARP per (image, cell).  The reference calls cv::FAST(cell, iniTh) and, if that returns NO keypoint after
// non-max suppression, cv::FAST(cell, minTh) (:811-817).  Both calls are trips through the same passes with
//   m(p)      = max over the 16 arcs of 9 contiguous circle pixels of min |v - p_k| (same-sign arcs only)
//   corner@t  <=> m > t,  cornerScore = m - 1  (independent of t for a corner)
// and NMS as a strict '>' against the 8 neighbours (non-corners hold 0, the cell's outer ring holds 0: NMS is
// cell-local).  Rounds 1 v1-v5 folded the two calls into ONE trip at min(iniTh, minTh) (keep@t = keep@tlo && s >= t);
// v6 runs the iniTh trip first and the minTh trip only for cells it left empty: a textured cell lists 3-4x fewer
// pixels for the exact arc test at 20 than at 7, which outweighs the second prefilter pass of low-contrast cells
// (1.24 -> 1.09 ms per 512 images).  Survivors leave in row-major order (the reference's emission order) via a
// per-row 64-bit bitmap + popcount scan -- no atomics-ordered output.
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
//
// Round 2, late: grid = (cell groups, images) instead of one linear index split by a 64-bit division; both halves of the row
// count scan in one register; pass 1a branch-free up to the ballot with the column limits moved to pass 1b and the word
// list holding tile word indices (57 -> 47 instructions per 32 words): 1.075 -> 1.039 ms per 512 images.  ncu: issue slots
// 80 % busy, ALU pipe 70 %, math-pipe throttle the first stall reason -- what is left is instruction count on the ALU pipe.
//
// Open (evaluated, not built; neither changes results):
//  * band tiles: one warp takes a vertical band of cells and keeps the three rows above / below in registers across
//    the cell boundary, so the +-3-row words of pass (1) and the ring rows of pass (3) are loaded once per band instead
//    of once per cell; estimated -15 % of this kernel (-5 % of the step), at the price of per-band instead of per-cell
//    threshold retry bookkeeping.
//  * the minTh trip's prefilter flags could be produced during the iniTh trip (same loads, second SWAR compare) and
//    kept as a row bitmap for the cells that turn out empty: removes the second pass (1) for ~half of the cells,
//    adds ~28 % to pass (1) of every cell: net -6 % of this kernel; needs 512 B more shared memory per warp, which
//    costs one resident CTA per SM.
#include <type_traits>

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
    int img0, nimg, cells_total, cell_cap, t_ini, t_min;
    int tile_bytes, sc_bytes, plist_bytes, warp_bytes;    // per-warp shared layout, see fast_smem_layout
    const uint4* cell_tab;
    uint32_t* cell_cnt; uint32_t* cell_pts; int32_t* err;
};

namespace {
#ifndef GFORB_FAST_WARPS
#define GFORB_FAST_WARPS 4      // measured: 2 warps 1.048 ms, 4 warps 1.040, 8 warps 1.064
#endif
constexpr int kWarpsPerCta = GFORB_FAST_WARPS;
constexpr int kBitmapBytes = 64 * 2 * 4;         // keep@tlo, 64 rows x 2 x 32 bits
// Per-warp shared layout: [tile (TMA, 128-byte aligned)][score map][row bitmap 512 B][pixel list][mbarrier 8 B],
// regions rounded to 16 bytes, the whole to 128.  The score map covers the FAST domain plus a one-pixel zero ring
// ((dh + 2) rows of align4(dw + 2) bytes), not the whole tile.  The word list of pass 1 is dead before the score map
// is first written, so it lives there (a cell has <= dh * (dw / 4 + 2) words: always fits).  At 752x480 this is
// 7,040 B per warp: 8 CTAs x 4 warps per SM.
// per-byte "x != 0" for words whose bytes have bit 0 clear: bit 7 of each lane
__device__ __forceinline__ uint32_t nz_lanes(uint32_t x) { return (x >> 1) + 0x7f7f7f7fu; }
}  // namespace

// One WARP per (image, cell): no block barriers, no block-wide atomics; the four warps of a CTA are independent.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
fast_cells_kernel(const FastArgs a, const __grid_constant__ TileMaps maps) {
    extern __shared__ __align__(128) uint8_t smem[];

    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int cell = blockIdx.x * kWarpsPerCta + wid;       // grid = (cell groups, images): no division to split a linear index
    if (cell >= a.cells_total) return;
    const int img = a.img0 + blockIdx.y;
    const size_t out_cell = (size_t)img * a.cells_total + cell;
    const uint4 ca = __ldg(a.cell_tab + 2 * cell), cb = __ldg(a.cell_tab + 2 * cell + 1);
    const int cw = (ca.y >> 8) & 0xFF, ch = (ca.y >> 16) & 0xFF;
    if (cw == 0) {                                      // cell skipped by the reference (:796,805) or degenerate
        if (lane == 0) a.cell_cnt[out_cell] = 0;
        return;
    }
    uint8_t* wbase = smem + (size_t)wid * a.warp_bytes;
    uint8_t* tile = wbase;
    uint8_t* sc = wbase + a.tile_bytes;                                   // corner scores: domain + 1-px zero ring, pitch sp
    uint32_t* keep_lo = reinterpret_cast<uint32_t*>(sc + a.sc_bytes);     // [row][2]: survivors of NMS at threshold tlo
    uint16_t* list = reinterpret_cast<uint16_t*>(keep_lo + 128);          // (row << 8) | tile column
    uint16_t* wlist = reinterpret_cast<uint16_t*>(sc);                    // pass 1 only (before the score map is cleared)
    uint64_t* bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(list) + a.plist_bytes);

    const int xoff = ca.y & 0xFF, l = ca.y >> 24;
    const int tp = ca.w & 0xFF, th = (ca.w >> 8) & 0xFF, nw = ca.w >> 24;
    const int dw = cw - 6, dh = ch - 6;                 // FAST domain: rows/cols 3 .. dim-4 of the cell image

    if (lane == 0) {
        mbar_init(bar, 1);
        mbar_expect_tx(bar, (uint32_t)(tp * th));
        tma_load_tile_3d(tile, &maps.fast[l], (int)(ca.x & 0xFFFF), (int)(ca.x >> 16), img, bar);
    }
    const int sp = (dw + 2 + 3) & ~3;                   // score map pitch
    reinterpret_cast<uint4*>(keep_lo)[lane] = make_uint4(0, 0, 0, 0);     // row bitmaps
    const int t_ini = min(max(a.t_ini, 0), 255), t_min = min(max(a.t_min, 0), 255);
    const int tpw = tp >> 2;
    const int xlo = xoff + 3, xhi = xlo + dw;           // domain columns [xlo, xhi) in tile bytes
    const int w0 = xlo >> 2;
    uint8_t* scb = sc - 2 * sp - xlo + 1;               // scb[row * sp + col] = score of tile pixel (row, col)
    const int nitems = dh * nw;                         // item = one 32-bit word of a domain row
    const int qs = (int)(cb.y & 0xFF), rs = (int)(cb.y >> 8);         // 32 = qs * nw + rs
    const int ir0 = (lane * (int)cb.x) >> 16, iw0 = lane - ir0 * nw;  // cb.x = ceil(65536 / nw): lane -> (row, word)
    const uint32_t c4_first = (0xFu << (xlo & 3)) & 0xFu;            // pixels of the first / last word column inside [xlo, xhi)
    const uint32_t c4_last = 0xFu >> (3 - ((xhi - 1) & 3));
    const uint32_t inv_tpw = cb.z;                                     // ceil(65536 / tpw)
    const uint32_t lt = (1u << lane) - 1u;
    const int pstep = qs * tpw + rs, pwrap = tpw - nw;
    __syncwarp();                                       // barrier init + clears visible to the whole warp
    mbar_wait(bar, 0);

    // The reference's control flow, literally: cv::FAST(cell, iniThFAST); only if that leaves NO keypoint after NMS,
    // cv::FAST(cell, minThFAST) (src/ORBextractor.cc:811-817).  v6: the second threshold is a second trip through the
    // passes below instead of being folded into one trip at min(ini, min): a textured cell (about half of them) is done
    // after the cheap trip at 20 -- far fewer pixels survive the prefilter and reach the exact arc test than at 7 -- and
    // only the low-contrast cells pay for both.
    unsigned long long m0 = 0, m1 = 0;                  // NMS survivors of rows lane and lane + 32
    for (int trip = 0; trip < 2; ++trip) {
        const int tc = trip ? t_min : t_ini;
        // prefilter threshold: the largest 2^k - 1 <= tc (a lower threshold stays a superset); |d| > 2^k - 1 <=> d & ~(2^k - 1) != 0,
        // and (|d1| | |d2|) & mask != 0 <=> either exceeds it
        const uint32_t pmask = (~(uint32_t)((1 << (31 - __clz(min(tc, 127) + 1))) - 1) & 0xFFu) * 0x01010101u;

        // ---- pass 1a: 4-pixel prefilter, 32 words per step; words with any flagged pixel are compacted by ONE ballot
        //      into a word list (tile word index << 4 | 4-bit pixel mask).  Branch-free up to the ballot: a lane past the
        //      last item reads a safe word and drops its flags; the domain's column limits are applied in pass 1b, on the
        //      listed words only.  tc == 0 ("> 0" has no cleared-low-bit form) lists the whole domain: its own loop ----
        const uint32_t* const tile32 = reinterpret_cast<const uint32_t*>(tile);
        const uint32_t* const psafe = tile32 + 3 * tpw + 1;                 // rows 0..6, words 0..2: inside every tile
        int nwl = 0;
        auto prefilter = [&](auto all_tag) {
            constexpr bool kAll = decltype(all_tag)::value;
            int iw = iw0;
            const uint32_t* p = tile32 + (3 + ir0) * tpw + w0 + iw0;
            for (int i0 = 0; i0 < nitems; i0 += 32) {
                const bool ok = i0 + lane < nitems;
                uint32_t f;
                if (kAll) f = ok ? 0x80808080u : 0u;
                else {
                    const uint32_t* q = ok ? p : psafe;
                    const uint32_t C = q[0], Wm = q[-1], Wp = q[1], U = q[-3 * tpw], D = q[3 * tpw];
                    const uint32_t Lw = __funnelshift_r(Wm, C, 8), Rw = __funnelshift_r(C, Wp, 24);
                    const uint32_t v = (__vabsdiffu4(C, U) | __vabsdiffu4(C, D)) & pmask;
                    const uint32_t h = (__vabsdiffu4(C, Lw) | __vabsdiffu4(C, Rw)) & pmask;
                    f = ok ? (nz_lanes(v) & nz_lanes(h) & 0x80808080u) : 0u;
                }
                const uint32_t m = __ballot_sync(0xffffffffu, f != 0);
                if (f != 0) {
                    // flag bits 7, 15, 23, 31 -> mask bits 0..3: the four 0/1 bytes times 0x01020408 meet in the top byte
                    const uint32_t m4 = (((f >> 7) & 0x01010101u) * 0x01020408u) >> 24;
                    wlist[nwl + __popc(m & lt)] = (uint16_t)(((uint32_t)(p - tile32) << 4) | m4);
                }
                nwl += __popc(m);
                iw += rs; p += pstep;
                if (iw >= nw) { iw -= nw; p += pwrap; }
            }
        };
        if (tc == 0) prefilter(std::true_type{});
        else prefilter(std::false_type{});
        __syncwarp();
        // ---- pass 1b: expand the word list into the pixel list ((row << 8) | tile column); order is irrelevant.
        //      Columns outside [xlo, xhi) leave here (first / last word column of the domain).
        //      Per-lane counts are <= 4: three ballots give the exclusive prefix ----
        int nlist = 0;
        for (int k0 = 0; k0 < nwl; k0 += 32) {
            const uint32_t we = k0 + lane < nwl ? wlist[k0 + lane] : 0u;
            const uint32_t widx = we >> 4;
            const uint32_t row = (widx * inv_tpw) >> 16, wc = widx - row * (uint32_t)tpw;     // exact: widx < 66 * 64
            uint32_t m4 = we & 15u;
            if ((int)wc == w0) m4 &= c4_first;
            if ((int)wc == w0 + nw - 1) m4 &= c4_last;
            const int c = __popc(m4);
            const uint32_t b0 = __ballot_sync(0xffffffffu, c & 1), b1 = __ballot_sync(0xffffffffu, c & 2);
            const uint32_t b2 = __ballot_sync(0xffffffffu, c & 4);
            uint16_t* out = list + nlist + __popc(b0 & lt) + 2 * __popc(b1 & lt) + 4 * __popc(b2 & lt);
            nlist += __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
            const uint32_t e = (row << 8) | (wc << 2);
            if (m4 & 1u) *out++ = (uint16_t)e;
            if (m4 & 2u) *out++ = (uint16_t)(e + 1);
            if (m4 & 4u) *out++ = (uint16_t)(e + 2);
            if (m4 & 8u) *out++ = (uint16_t)(e + 3);
        }
        __syncwarp();
        {   // the word list is dead: clear the score map (domain + zero ring) it shared
            const int n16 = ((dh + 2) * sp + 15) >> 4;
            for (int i = lane; i < n16; i += 32) reinterpret_cast<uint4*>(sc)[i] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();

        // ---- pass 2: exact corner strength of the listed pixels; corner <=> strength > tc, score = strength - 1.
        //      Corners are compacted in place to the front of the list (ballot), so that NMS visits corners only ----
        int ncorner = 0;
        for (int k0 = 0; k0 < nlist; k0 += 32) {
            const int k = k0 + lane;
            int e = 0, m = 0;
            if (k < nlist) {
                e = list[k];
                const uint8_t* c = tile + (e >> 8) * tp + (e & 0xFF);
                const uint8_t* const rows[7] = {c - 3 * tp, c - 2 * tp, c - tp, c, c + tp, c + 2 * tp, c + 3 * tp};
                m = fast_arc_strength(rows, c[0]);
            }
            const bool corner = k < nlist && m > tc;
            const uint32_t cm = __ballot_sync(0xffffffffu, corner);
            __syncwarp();                                // every lane has read its entry before the slots are reused
            if (corner) {
                scb[(e >> 8) * sp + (e & 0xFF)] = (uint8_t)(m - 1);
                list[ncorner + __popc(cm & lt)] = (uint16_t)e;
            }
            ncorner += __popc(cm);
        }
        __syncwarp();

        // ---- pass 3: strict 3x3 NMS on the score map (zero outside the domain) over the corners ----
        for (int k = lane; k < ncorner; k += 32) {
            const int e = list[k];
            const int py = e >> 8, px = e & 0xFF;
            const uint8_t* q = scb + py * sp + px;
            const int s = q[0];
            const int n0 = max(max(q[-sp - 1], q[-sp]), q[-sp + 1]), n1 = max(q[-1], q[1]);
            const int n2 = max(max(q[sp - 1], q[sp]), q[sp + 1]);
            if (s > max(max(n0, n1), n2)) {
                const int bx = px - xlo;
                atomicOr(&keep_lo[2 * (py - 3) + (bx >> 5)], 1u << (bx & 31));
            }
        }
        __syncwarp();
        if (lane < dh) m0 = (unsigned long long)keep_lo[2 * lane] | ((unsigned long long)keep_lo[2 * lane + 1] << 32);
        if (lane + 32 < dh) m1 = (unsigned long long)keep_lo[2 * lane + 64] | ((unsigned long long)keep_lo[2 * lane + 65] << 32);
        if (__any_sync(0xffffffffu, (m0 | m1) != 0) || t_min == t_ini) break;     // keypoints found: no retry (:813)
    }
    // exclusive scan of the <= 59 row counts; both halves ride in one register (a half's total is <= 32 * 59 < 65536)
    const int c0 = __popcll(m0), c1 = __popcll(m1);
    int i01 = c0 | (c1 << 16);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, i01, o);
        if (lane >= o) i01 += t;
    }
    const int tot01 = __shfl_sync(0xffffffffu, i01, 31);
    const int i0 = i01 & 0xFFFF, i1 = i01 >> 16, tot0 = tot01 & 0xFFFF, tot1 = tot01 >> 16;
    int total = tot0 + tot1;
    if (total > a.cell_cap) {
        if (lane == 0) atomicOr(a.err, kErrCellOverflow);
        total = 0;
    }
    if (lane == 0) a.cell_cnt[out_cell] = (uint32_t)total;
    if (total == 0) return;
    const int ox = (int)(ca.z & 0xFFFF), oy = (int)(ca.z >> 16);          // j*wCell, i*hCell (:824-825)
    uint32_t* out0 = a.cell_pts + out_cell * (size_t)a.cell_cap;
    {
        uint32_t* out = out0 + (i0 - c0);
        const int yy = lane + 3;
        for (unsigned long long r = m0; r; r &= r - 1) {
            const int bx = __ffsll((long long)r) - 1;
            *out++ = pack_cand(ox + bx + 3, oy + yy, scb[yy * sp + bx + xlo]);
        }
    }
    if (c1) {
        uint32_t* out = out0 + tot0 + (i1 - c1);
        const int yy = lane + 35;
        for (unsigned long long r = m1; r; r &= r - 1) {
            const int bx = __ffsll((long long)r) - 1;
            *out++ = pack_cand(ox + bx + 3, oy + yy, scb[yy * sp + bx + xlo]);
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
            const int nw = ((xhi - 1) >> 2) - (xlo >> 2) + 1;           // <= 16 word columns
            const int R = 0;     // (unused since the warp-per-cell kernel)
            e[0] = (uint32_t)(gx - xoff) | ((uint32_t)(kEdge + iniY) << 16);
            e[1] = (uint32_t)xoff | ((uint32_t)cw << 8) | ((uint32_t)ch << 16) | ((uint32_t)l << 24);
            e[2] = (uint32_t)(cj * L.wCell) | ((uint32_t)(ci * L.hCell) << 16);
            e[3] = (uint32_t)L.fast_bw | ((uint32_t)L.fast_bh << 8) | ((uint32_t)R << 16) | ((uint32_t)nw << 24);
            e[4] = (uint32_t)((65536 + nw - 1) / nw);
            e[5] = (uint32_t)(32 / nw) | ((uint32_t)(32 % nw) << 8);
            e[6] = (uint32_t)((65536 + L.fast_bw / 4 - 1) / (L.fast_bw / 4));
        }
    }
}

namespace {
struct FastLayout { int tile_bytes, sc_bytes, plist_bytes, warp_bytes; };
FastLayout fast_smem_layout(const Geom& g) {
    auto up = [](int v, int a) { return (v + a - 1) / a * a; };
    int tile = 128, scm = 128, pl = 64;
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        if (L.ncells == 0) continue;
        tile = std::max(tile, L.fast_bw * L.fast_bh);
        const int dwm = L.wCell, dhm = L.hCell;                       // largest FAST domain of the level
        const int words = dhm * ((dwm + 3) / 4 + 1);                  // a row of dw pixels spans <= ceil(dw / 4) + 1 words
        scm = std::max(scm, std::max((dhm + 2) * ((dwm + 2 + 3) & ~3) + 16, 2 * words));
        pl = std::max(pl, dwm * dhm);
    }
    FastLayout f;
    f.tile_bytes = up(tile, 16); f.sc_bytes = up(scm, 16); f.plist_bytes = up(2 * pl, 16);
    f.warp_bytes = up(f.tile_bytes + f.sc_bytes + kBitmapBytes + f.plist_bytes + 8, 128);
    return f;
}
size_t fast_cta_smem(const FastLayout& f) { return (size_t)kWarpsPerCta * f.warp_bytes; }
}  // namespace

int fast_configure(const Geom& g) {
    static size_t have[64] = {};       // high-water mark: engines of different footprints share the kernel symbol
    return opt_in_smem(fast_cells_kernel, fast_cta_smem(fast_smem_layout(g)), have) ? (int)cudaErrorInvalidValue : 0;
}

void launch_fast(const Geom& g, const TileMaps& m, const Buffers& b, int nimg, cudaStream_t s) {
    FastArgs a;
    a.img0 = g.img0; a.nimg = nimg; a.cells_total = g.cells_total; a.cell_cap = g.cell_cap;
    a.t_ini = g.ini_th; a.t_min = g.min_th;
    const FastLayout f = fast_smem_layout(g);
    a.tile_bytes = f.tile_bytes; a.sc_bytes = f.sc_bytes; a.plist_bytes = f.plist_bytes; a.warp_bytes = f.warp_bytes;
    a.cell_tab = reinterpret_cast<const uint4*>(b.cell_tab);
    a.cell_cnt = b.cell_cnt; a.cell_pts = b.cell_pts; a.err = b.err;
    if (nimg <= 0 || g.cells_total == 0) return;
    const dim3 grd((unsigned)((g.cells_total + kWarpsPerCta - 1) / kWarpsPerCta), (unsigned)nimg);
    fast_cells_kernel<<<grd, kWarpsPerCta * 32, fast_cta_smem(f), s>>>(a, m);
}

}  // namespace gforb
