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
// Compute: (1) every domain pixel takes the 4-pixel high-speed test at tlo, survivors are compacted into a
// shared list with warp ballots; (2) list entries get the exact arc strength from a 3-input packed-s16x2
// min/max network (VIMNMX3: both polarities in one register); (3) NMS walks the same list.
// Integer/LSU-bound: algorithmic bytes = level pixels read once + 4 B per survivor.
#include "gforb_internal.cuh"
#include "tma.cuh"

namespace gforb {

namespace {
constexpr int kMaxTP = 80;                // largest tile pitch: align16(59 + 6)
constexpr int kTileRows = kMaxCellDim;    // 66
constexpr int kThreads = 128;
constexpr int kMaxDomain = 59 * 59;

// Bresenham circle r=3 (src/FAST_NEON.cc:3-7); tp = tile pitch
__device__ __forceinline__ int circle_off(int k, int tp) {
    constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return dy[k] * tp + dx[k];
}

// max over the 16 arcs of min over 9 contiguous, for d = v-p (low half) and -d (high half) at once
__device__ __forceinline__ int fast_arc_strength(const uint8_t* c, int v, int tp) {
    uint32_t p[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int q = (int)c[circle_off(k, tp)];
        p[k] = __byte_perm((uint32_t)(v - q), (uint32_t)(q - v), 0x5410);   // {lo16(v-q), lo16(q-v)}
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
    return max((int)(short)(b0 & 0xFFFFu), (int)(short)(b0 >> 16));
}
}  // namespace

__global__ void __launch_bounds__(kThreads)
fast_cells_kernel(const Geom g, const __grid_constant__ TileMaps maps, uint32_t* __restrict__ cell_cnt,
                  uint32_t* __restrict__ cell_pts, int32_t* __restrict__ err) {
    __shared__ __align__(128) uint8_t tile_raw[kTileRows * kMaxTP];
    __shared__ __align__(16) uint8_t sc_raw[kTileRows * kMaxTP + 16];
    __shared__ uint16_t list[kMaxDomain + 3];          // (y << 8) | x of pixels passing the high-speed test
    __shared__ unsigned long long keep_lo[64], keep_hi[64];
    __shared__ int row_base[64];
    __shared__ __align__(8) uint64_t bar;
    __shared__ int s_any_hi, s_nlist;

    const int img = g.img0 + blockIdx.y;
    int cell = blockIdx.x;
    int l = 0;
    while (l + 1 < g.nlevels && cell >= g.lv[l + 1].cell_base) ++l;
    const LevelGeom& L = g.lv[l];
    cell -= L.cell_base;
    const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
    const size_t out_cell = (size_t)img * g.cells_total + blockIdx.x;

    const int iniX = kMinBorder + cj * L.wCell, iniY = kMinBorder + ci * L.hCell;
    // src/ORBextractor.cc:796,805: asymmetric skip tests, as written
    const int cw = min(iniX + L.wCell + 6, L.maxBX) - iniX;
    const int ch = min(iniY + L.hCell + 6, L.maxBY) - iniY;
    const int dw = cw - 6, dh = ch - 6;   // FAST domain: rows/cols 3 .. dim-4 of the cell image
    if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6 || dw <= 0 || dh <= 0) {
        if (threadIdx.x == 0) cell_cnt[out_cell] = 0;
        return;
    }
    const int tp = L.fast_bw;
    const int gx = kXPad + iniX, xoff = gx & 15;
    const uint8_t* tile = tile_raw + xoff;    // cell pixel (0,0)
    uint8_t* sc = sc_raw + xoff;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        s_any_hi = 0; s_nlist = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, (uint32_t)(tp * L.fast_bh));
        tma_load_tile_3d(tile_raw, &maps.fast[l], gx - xoff, kEdge + iniY, img, &bar);
    }
    // while the tile is in flight: clear the score map rows we will touch and the row bitmaps
    for (int i = threadIdx.x; i < (ch * tp + 16 + 3) / 4; i += kThreads) reinterpret_cast<uint32_t*>(sc_raw)[i] = 0;
    if (threadIdx.x < 64) { keep_lo[threadIdx.x] = 0; keep_hi[threadIdx.x] = 0; }

    const int t_ini = min(max(g.ini_th, 0), 255), t_min = min(max(g.min_th, 0), 255);
    const int tlo = min(t_ini, t_min);
    const int npx = dw * dh;
    const int step_y = kThreads / dw, step_x = kThreads - step_y * dw;
    int y = (int)threadIdx.x / dw, x = (int)threadIdx.x - y * dw;
    const int lane = threadIdx.x & 31;

    mbar_wait(&bar, 0);

    // pass 1: high-speed test (a 9-arc covers one pixel of each opposite pair) + ballot compaction
    for (int i0 = 0; i0 < npx; i0 += kThreads) {
        bool pass = false;
        if (i0 + (int)threadIdx.x < npx) {
            const uint8_t* c = tile + (y + 3) * tp + x + 3;
            const int v = c[0];
            const int d0 = v - c[3 * tp], d8 = v - c[-3 * tp], d4 = v - c[3], d12 = v - c[-3];
            const bool dark = (d0 > tlo || d8 > tlo) && (d4 > tlo || d12 > tlo);
            const bool bright = (d0 < -tlo || d8 < -tlo) && (d4 < -tlo || d12 < -tlo);
            pass = dark || bright;
        }
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(&s_nlist, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (pass) list[base + __popc(m & ((1u << lane) - 1))] = (uint16_t)(((y + 3) << 8) | (x + 3));
        }
        x += step_x; y += step_y;
        if (x >= dw) { x -= dw; ++y; }
    }
    __syncthreads();
    const int nlist = s_nlist;

    // pass 2: exact corner strength of the listed pixels
    for (int k = threadIdx.x; k < nlist; k += kThreads) {
        const int e = list[k];
        const int off = (e >> 8) * tp + (e & 0xFF);
        const int m = fast_arc_strength(tile + off, tile[off], tp);
        if (m > tlo) sc[off] = (uint8_t)(m - 1);
    }
    __syncthreads();

    // pass 3: strict 3x3 NMS on the score map (zero outside the domain), again only over the list
    for (int k = threadIdx.x; k < nlist; k += kThreads) {
        const int e = list[k];
        const int py = e >> 8, px = e & 0xFF;
        const uint8_t* q = sc + py * tp + px;
        const int s = q[0];
        if (s == 0) continue;
        const bool keep = s > q[-1] && s > q[1] && s > q[-tp - 1] && s > q[-tp] && s > q[-tp + 1] &&
                          s > q[tp - 1] && s > q[tp] && s > q[tp + 1];
        if (!keep) continue;
        const unsigned long long bit = 1ull << (px - 3);
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
                    if (sc[(threadIdx.x + 3) * tp + bx + 3] >= t_min) f |= 1ull << bx;
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
            if (total > g.cell_cap) { atomicOr(err, kErrCellOverflow); total = 0; }
            cell_cnt[out_cell] = (uint32_t)total;
            s_nlist = total;
        }
    }
    __syncthreads();
    if (s_nlist == 0) return;
    if ((int)threadIdx.x < dh) {
        uint32_t* out = cell_pts + out_cell * (size_t)g.cell_cap + row_base[threadIdx.x];
        const int yy = threadIdx.x + 3;
        for (unsigned long long r = keep_lo[threadIdx.x]; r; r &= r - 1) {
            const int xx = __ffsll((long long)r) - 1 + 3;
            *out++ = pack_cand(cj * L.wCell + xx, ci * L.hCell + yy, sc[yy * tp + xx]);
        }
    }
}

void launch_fast(const Geom& g, const TileMaps& m, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd(g.cells_total, nimg);
    fast_cells_kernel<<<grd, kThreads, 0, s>>>(g, m, b.cell_cnt, b.cell_pts, b.err);
}

}  // namespace gforb
