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
// Integer/LSU-bound: algorithmic bytes = level pixels read once + 4 B per survivor.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kTP = 72;                   // smem tile pitch
constexpr int kTileRows = kMaxCellDim;    // 66
constexpr int kThreads = 128;

// Bresenham circle r=3 (src/FAST_NEON.cc:3-7)
__device__ __forceinline__ int circle_off(int k) {
    constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return dy[k] * kTP + dx[k];
}

// max over arcs of min over 9 contiguous, for d (low half) and -d (high half) at once
__device__ __forceinline__ int fast_arc_strength(const uint8_t* c, int v) {
    uint32_t p[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int d = v - (int)c[circle_off(k)];
        p[k] = ((uint32_t)d & 0xFFFFu) | ((uint32_t)(-d) << 16);
    }
    uint32_t m2[16], m4[16], m8[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m2[k] = __vmins2(p[k], p[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m4[k] = __vmins2(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; ++k) m8[k] = __vmins2(m4[k], m4[(k + 4) & 15]);
    uint32_t best = __vmins2(m8[0], p[8]);
#pragma unroll
    for (int k = 1; k < 16; ++k) best = __vmaxs2(best, __vmins2(m8[k], p[(k + 8) & 15]));
    const int a = (short)(best & 0xFFFFu), b = (short)(best >> 16);
    return max(a, b);
}
}  // namespace

__global__ void __launch_bounds__(kThreads)
fast_cells_kernel(const Geom g, const uint8_t* __restrict__ pyr, uint32_t* __restrict__ cell_cnt,
                  uint32_t* __restrict__ cell_pts, int32_t* __restrict__ err) {
    __shared__ __align__(16) uint8_t tile[kTileRows * kTP];
    __shared__ __align__(16) uint8_t sc[kTileRows * kTP];
    __shared__ unsigned long long keep_lo[64], keep_hi[64];
    __shared__ int row_base[64];
    __shared__ int s_any_hi;

    const int img = blockIdx.y;
    int cell = blockIdx.x;
    int l = 0;
    while (l + 1 < g.nlevels && cell >= g.lv[l + 1].cell_base) ++l;
    const LevelGeom& L = g.lv[l];
    cell -= L.cell_base;
    const int ci = cell / L.nCols, cj = cell - ci * L.nCols;
    const size_t out_cell = (size_t)img * g.cells_total + blockIdx.x;

    const int iniX = kMinBorder + cj * L.wCell, iniY = kMinBorder + ci * L.hCell;
    // src/ORBextractor.cc:796,805: asymmetric skip tests, as written
    if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6) {
        if (threadIdx.x == 0) cell_cnt[out_cell] = 0;
        return;
    }
    const int cw = min(iniX + L.wCell + 6, L.maxBX) - iniX;
    const int ch = min(iniY + L.hCell + 6, L.maxBY) - iniY;
    const int dw = cw - 6, dh = ch - 6;   // FAST domain: rows/cols 3 .. dim-4 of the cell image
    if (dw <= 0 || dh <= 0) {
        if (threadIdx.x == 0) cell_cnt[out_cell] = 0;
        return;
    }

    const uint8_t* src = pyr + L.off + (size_t)img * L.img_stride + (size_t)(kEdge + iniY) * L.pitch + kXPad + iniX;
    for (int i = threadIdx.x; i < ch * cw; i += kThreads) {
        const int y = i / cw, x = i - y * cw;
        tile[y * kTP + x] = __ldg(src + (size_t)y * L.pitch + x);
    }
    for (int i = threadIdx.x; i < kTileRows * kTP / 4; i += kThreads) reinterpret_cast<uint32_t*>(sc)[i] = 0;
    if (threadIdx.x < 64) { keep_lo[threadIdx.x] = 0; keep_hi[threadIdx.x] = 0; }
    if (threadIdx.x == 0) s_any_hi = 0;
    __syncthreads();

    const int t_ini = min(max(g.ini_th, 0), 255), t_min = min(max(g.min_th, 0), 255);
    const int tlo = min(t_ini, t_min);

    // pass 1: corner strength at the lower threshold
    for (int i = threadIdx.x; i < dw * dh; i += kThreads) {
        const int y = i / dw + 3, x = i - (y - 3) * dw + 3;
        const uint8_t* c = tile + y * kTP + x;
        const int v = c[0];
        // high-speed reject: a 9-arc covers one of each opposite pair
        const int d0 = v - c[circle_off(0)], d8 = v - c[circle_off(8)];
        const int d4 = v - c[circle_off(4)], d12 = v - c[circle_off(12)];
        const bool dark = (d0 > tlo || d8 > tlo) && (d4 > tlo || d12 > tlo);
        const bool bright = (d0 < -tlo || d8 < -tlo) && (d4 < -tlo || d12 < -tlo);
        if (!(dark || bright)) continue;
        const int m = fast_arc_strength(c, v);
        if (m > tlo) sc[y * kTP + x] = (uint8_t)(m - 1);
    }
    __syncthreads();

    // pass 2: strict 3x3 NMS on the score map (zero outside the domain)
    for (int i = threadIdx.x; i < dw * dh; i += kThreads) {
        const int y = i / dw + 3, x = i - (y - 3) * dw + 3;
        const uint8_t* q = sc + y * kTP + x;
        const int s = q[0];
        if (s == 0) continue;
        const bool keep = s > q[-1] && s > q[1] && s > q[-kTP - 1] && s > q[-kTP] && s > q[-kTP + 1] &&
                          s > q[kTP - 1] && s > q[kTP] && s > q[kTP + 1];
        if (!keep) continue;
        const unsigned long long bit = 1ull << (x - 3);
        atomicOr(&keep_lo[y - 3], bit);
        if (s >= t_ini) { atomicOr(&keep_hi[y - 3], bit); s_any_hi = 1; }
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
                    if (sc[(threadIdx.x + 3) * kTP + bx + 3] >= t_min) f |= 1ull << bx;
                }
                m = f;
            }
        }
        keep_lo[threadIdx.x] = m;
        row_base[threadIdx.x] = __popcll(m);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int r = 0; r < dh; ++r) { const int c = row_base[r]; row_base[r] = acc; acc += c; }
        if (acc > g.cell_cap) { atomicOr(err, kErrCellOverflow); acc = 0; }
        cell_cnt[out_cell] = (uint32_t)acc;
        s_any_hi = acc;
    }
    __syncthreads();
    if (s_any_hi == 0) return;
    if ((int)threadIdx.x < dh) {
        uint32_t* out = cell_pts + out_cell * (size_t)g.cell_cap + row_base[threadIdx.x];
        const int y = threadIdx.x + 3;
        for (unsigned long long r = keep_lo[threadIdx.x]; r; r &= r - 1) {
            const int x = __ffsll((long long)r) - 1 + 3;
            *out++ = pack_cand(cj * L.wCell + x, ci * L.hCell + y, sc[y * kTP + x]);
        }
    }
}

void launch_fast(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd(g.cells_total, nimg);
    fast_cells_kernel<<<grd, kThreads, 0, s>>>(g, b.pyr, b.cell_cnt, b.cell_pts, b.err);
}

}  // namespace gforb
