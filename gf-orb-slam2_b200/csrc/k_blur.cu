// 7x7 sigma=2 Gaussian blur of every pyramid level: the cv::GaussianBlur call of ORBextractor::operator()
// (reference src/ORBextractor.cc:1153-1155) in OpenCV's 8-bit fixed-point form: Q8 kernel
// {18,34,48,56,48,34,18} (sum 256), rows accumulated exactly in 16 bits, columns in 32 bits,
// result (v + 32768) >> 16.  BORDER_REFLECT_101 comes for free: the source is the padded pyramid whose
// 19-px frame IS the reflection of the ROI (the reference blurs a clone of the ROI, i.e. reflects from the ROI).
//
// Register-resident separable filter, no shared memory: a warp owns a 128-column strip of <= 56 rows (balanced per
// level), a lane 4 adjacent output columns.  Per input row a lane loads three aligned 32-bit words (12 bytes;
// coalesced across the warp), forms the four 7-tap row sums with funnel shifts + two IDP4A each, and packs them
// with the previous row's sums into 16-bit ROW PAIRS; the column pass is four IDP2A per pixel over a ring of six
// live pairs (the row loop is unrolled by 6 so the ring rotates by renaming), then rounding and one packed 32-bit
// store.  DRAM latency is hidden by an L1 prefetch 8 rows ahead (CCTL.PF1) and by a straight-line 6-row body that
// lets the scheduler hoist each row's loads above the previous row's arithmetic.
// Streaming stencil, HBM-bound by design: algorithmic bytes = level pixels read + written (SURVEY.md 8d K5,
// "blur materialised" variant); ~13 instructions per pixel.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kWarps = 4;          // strips per CTA (stacked in y)
constexpr int kAhead = 8;          // rows of L1 prefetch distance
constexpr unsigned kK0123 = 18u | (34u << 8) | (48u << 16) | (56u << 24);
constexpr unsigned kK456 = 48u | (34u << 8) | (18u << 16);

struct Row3 { uint32_t a, b, c; };   // the three aligned words around a lane's 4 output columns
__device__ __forceinline__ Row3 load_row(const uint32_t* wp) { return Row3{__ldg(wp - 1), __ldg(wp), __ldg(wp + 1)}; }
__device__ __forceinline__ void row_sums(const Row3& r, int (&o)[4]) {
    o[0] = (int)__dp4a(__funnelshift_r(r.b, r.c, 8), kK456, __dp4a(__funnelshift_r(r.a, r.b, 8), kK0123, 0u));
    o[1] = (int)__dp4a(__funnelshift_r(r.b, r.c, 16), kK456, __dp4a(__funnelshift_r(r.a, r.b, 16), kK0123, 0u));
    o[2] = (int)__dp4a(__funnelshift_r(r.b, r.c, 24), kK456, __dp4a(__funnelshift_r(r.a, r.b, 24), kK0123, 0u));
    o[3] = (int)__dp4a(r.c, kK456, __dp4a(r.b, kK0123, 0u));
}
}  // namespace

__global__ void __launch_bounds__(kWarps * 32)
blur_kernel(const Geom g, const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur) {
    const int img = g.img0 + blockIdx.y;
    const int lane = threadIdx.x & 31;
    int t = blockIdx.x * kWarps + (threadIdx.x >> 5), l = 0;      // warp task = (level, strip, 128-column tile)
    if (t >= g.btiles_total) return;
    while (l + 1 < g.nlevels && t >= g.lv[l + 1].btile_base) ++l;
    const LevelGeom& L = g.lv[l];
    t -= L.btile_base;
    const int ts = t / L.btiles_x, tx = t - ts * L.btiles_x;
    const int X = tx * 128 + lane * 4;                       // first of this lane's 4 output columns
    const int y0 = ts * L.bstrip_rows;
    if (X >= L.w) return;
    const int y1 = min(y0 + L.bstrip_rows, L.h);
    const int nrows = y1 - y0;

    int pw = L.pitch >> 2, bpw = L.bpitch >> 2;              // pitches in words
    asm volatile("" : "+r"(pw), "+r"(bpw));                 // keep them in registers (else re-read from the parameter bank per row)
    // byte X of ROI row r: the ROI starts 32 bytes into a 64-byte-aligned padded row, X % 4 == 0 -> aligned words.
    // sp walks the input rows (starting 3 above the strip), dp the output rows.
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(
        pyr + L.off + (size_t)img * L.img_stride + (size_t)(kEdge + y0 - 3) * L.pitch + kXPad + X);
    uint32_t* dp = reinterpret_cast<uint32_t*>(blur + L.boff + (size_t)img * L.bimg_stride + (size_t)y0 * L.bpitch + X);

    // Vertical pass on ROW PAIRS: P_j = (H_j, H_j+1) packs the 16-bit row sums of input rows j, j+1 (same column), so
    // IDP2A applies two of the seven column taps per instruction: out_i = 32768 + P_i.(18,34) + P_i+2.(48,56) +
    // P_i+4.(48,34) + P_i+5.(0,18), rows counted from 3 above the strip.  Six live pairs per column, slot = j % 6
    // (the row loop is unrolled by 6 so the ring rotates by renaming).
    constexpr unsigned kW01 = 18u | (34u << 8), kW23 = 48u | (56u << 8), kW45 = 48u | (34u << 8), kW6 = 18u << 8;
    uint32_t P[6][4];
    int prevH[4], h[4];
    row_sums(load_row(sp), prevH);
    sp += pw;
#pragma unroll
    for (int j = 0; j < 5; ++j, sp += pw) {
        row_sums(load_row(sp), h);
#pragma unroll
        for (int c = 0; c < 4; ++c) { P[j][c] = __byte_perm((uint32_t)prevH[c], (uint32_t)h[c], 0x5410); prevH[c] = h[c]; }
    }
    // software pipeline: the words of the NEXT input row are requested before the current one is consumed.
    // The last strip row's prefetch reads one row further down: still inside the 19-px bottom frame.
    Row3 nxt = load_row(sp);
    // DRAM latency is far longer than one row of arithmetic: every row also pulls the row kAhead below it into L1
    // (lane 0 / 31 fetch the sector of their left / right neighbour word, the others their own)
    const uint32_t* pf = sp + (lane == 31) - (lane == 0);
#pragma unroll
    for (int j = 0; j < kAhead; ++j, pf += pw) asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
    const int pf_rows = L.h + kEdge - (y0 + 3) - kAhead;      // prefetches that stay inside the padded level

    // (Measured, round 2: a second copy of the body without the per-row `live` / prefetch-range tests for the groups of six
    // rows that lie wholly inside the strip -- 17 % fewer instructions in those groups -- ran at 0.367 ms against 0.358 ms
    // with the register count held at 64, 0.434 ms at the 80 registers the compiler wanted: rejected.)
    // The 6-row body is straight-line code (one basic block): rows past the end of the strip keep re-reading the last
    // input row (pointer stops advancing) and only their store is predicated off, so the scheduler can hoist every
    // row's loads above the previous row's arithmetic.
    for (int i0 = 0; i0 < nrows; i0 += 6) {
#pragma unroll
        for (int u = 0; u < 6; ++u) {
            const bool live = i0 + u < nrows;
            const Row3 cur = nxt;
            sp += live ? pw : 0;
            nxt = load_row(sp);
            if (i0 + u < pf_rows) asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
            pf += pw;
            row_sums(cur, h);
            uint32_t v[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                P[(u + 5) % 6][c] = __byte_perm((uint32_t)prevH[c], (uint32_t)h[c], 0x5410);
                prevH[c] = h[c];
                v[c] = __dp2a_lo(P[(u + 5) % 6][c], kW6,
                       __dp2a_lo(P[(u + 4) % 6][c], kW45,
                       __dp2a_lo(P[(u + 2) % 6][c], kW23,
                       __dp2a_lo(P[u][c], kW01, 32768u))));
            }
            // byte 2 of each sum is the pixel ((v + 32768) >> 16 <= 255)
            // bpitch is a multiple of 64: the word may spill past w into row padding, never past the pitch
            if (live) *dp = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            dp += bpw;
        }
    }
}

void launch_blur(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd((g.btiles_total + kWarps - 1) / kWarps, nimg);
    blur_kernel<<<grd, kWarps * 32, 0, s>>>(g, b.pyr, b.blur);
}

}  // namespace gforb
