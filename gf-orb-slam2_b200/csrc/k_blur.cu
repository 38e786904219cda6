// 7x7 sigma=2 Gaussian blur of every pyramid level: the cv::GaussianBlur call of ORBextractor::operator()
// (reference src/ORBextractor.cc:1153-1155) in OpenCV's 8-bit fixed-point form: Q8 kernel
// {18,34,48,56,48,34,18} (sum 256), rows accumulated exactly in 16 bits, columns in 32 bits,
// result (v + 32768) >> 16.  BORDER_REFLECT_101 comes for free: the source is the padded pyramid whose
// 19-px frame IS the reflection of the ROI (the reference blurs a clone of the ROI, i.e. reflects from the ROI).
//
// Register-resident separable filter, no shared memory: a lane owns 4 adjacent output columns and walks
// down a 32-row strip.  Per input row it loads three aligned 32-bit words (12 bytes; coalesced across the
// warp, 128 columns per warp), forms the four 7-tap row sums with funnel shifts + two IDP4A each, and
// pushes them into a 7-deep sliding window held in registers (the row loop is unrolled by 7 so the window
// rotates by renaming, not by moves); the column sum, rounding and a packed 32-bit store follow.
// Streaming stencil, HBM-bound by design: algorithmic bytes = level pixels read + written (SURVEY.md 8d K5,
// "blur materialised" variant); ~14 instructions per pixel.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kStripRows = 32;     // output rows per warp
constexpr int kWarps = 4;          // strips per CTA (stacked in y)
constexpr unsigned kK0123 = 18u | (34u << 8) | (48u << 16) | (56u << 24);
constexpr unsigned kK456 = 48u | (34u << 8) | (18u << 16);

__device__ __forceinline__ void row_sums(const uint8_t* row, int (&o)[4]) {
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(row);
    const uint32_t A = __ldg(wp - 1), B = __ldg(wp), C = __ldg(wp + 1);
    o[0] = (int)__dp4a(__funnelshift_r(B, C, 8), kK456, __dp4a(__funnelshift_r(A, B, 8), kK0123, 0u));
    o[1] = (int)__dp4a(__funnelshift_r(B, C, 16), kK456, __dp4a(__funnelshift_r(A, B, 16), kK0123, 0u));
    o[2] = (int)__dp4a(__funnelshift_r(B, C, 24), kK456, __dp4a(__funnelshift_r(A, B, 24), kK0123, 0u));
    o[3] = (int)__dp4a(C, kK456, __dp4a(B, kK0123, 0u));
}
}  // namespace

__global__ void __launch_bounds__(kWarps * 32)
blur_kernel(const Geom g, const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur) {
    const int img = g.img0 + blockIdx.y;
    int t = blockIdx.x, l = 0;
    while (l + 1 < g.nlevels && t >= g.lv[l + 1].btile_base) ++l;
    const LevelGeom& L = g.lv[l];
    t -= L.btile_base;
    const int ty = t / L.btiles_x, tx = t - ty * L.btiles_x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int X = tx * 128 + lane * 4;                       // first of this lane's 4 output columns
    const int y0 = (ty * kWarps + warp) * kStripRows;
    if (X >= L.w || y0 >= L.h) return;
    const int y1 = min(y0 + kStripRows, L.h);

    const int pitch = L.pitch;
    // byte X of ROI row r: the ROI starts 32 bytes into a 64-byte-aligned padded row, X % 4 == 0 -> aligned words
    const uint8_t* src = pyr + L.off + (size_t)img * L.img_stride + (size_t)kEdge * pitch + kXPad + X;
    uint8_t* dst = blur + L.boff + (size_t)img * L.bimg_stride + X;

    int win[7][4];
#pragma unroll
    for (int j = 0; j < 6; ++j) row_sums(src + (ptrdiff_t)(y0 - 3 + j) * pitch, win[j]);

    // input row r = y0 + 3 + i completes output row y0 + i; window slot of input row (y0-3+j) is j % 7
    for (int i0 = 0; i0 < kStripRows; i0 += 7) {
#pragma unroll
        for (int u = 0; u < 7; ++u) {
            const int y = y0 + i0 + u;
            if (y < y1) {
                // slots: newest row goes to (6 + u) % 7; centred taps are slots (u + k) % 7 for k = 0..6
                row_sums(src + (ptrdiff_t)(y + 3) * pitch, win[(6 + u) % 7]);
                uint32_t word = 0;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int v = 18 * (win[u % 7][c] + win[(u + 6) % 7][c]) + 34 * (win[(u + 1) % 7][c] + win[(u + 5) % 7][c]) +
                                  48 * (win[(u + 2) % 7][c] + win[(u + 4) % 7][c]) + 56 * win[(u + 3) % 7][c] + 32768;
                    word |= ((uint32_t)v >> 16) << (8 * c);
                }
                // bpitch is a multiple of 64: the word may spill past w into row padding, never past the pitch
                *reinterpret_cast<uint32_t*>(dst + (size_t)y * L.bpitch) = word;
            }
        }
    }
}

void launch_blur(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd(g.btiles_total, nimg);
    blur_kernel<<<grd, kWarps * 32, 0, s>>>(g, b.pyr, b.blur);
}

}  // namespace gforb
