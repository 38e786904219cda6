// 7x7 sigma=2 Gaussian blur of every pyramid level: the cv::GaussianBlur call of ORBextractor::operator()
// (reference src/ORBextractor.cc:1153-1155) in OpenCV's 8-bit fixed-point form: Q8 kernel
// {18,34,48,56,48,34,18} (sum 256), rows accumulated exactly in 16 bits, columns in 32 bits,
// result (v + 32768) >> 16.  BORDER_REFLECT_101 comes for free: the source is the padded pyramid whose
// 19-px frame IS the reflection of the ROI (the reference blurs a clone of the ROI, i.e. reflects from the ROI).
//
// One CTA = one 64x16 output tile of one level of one image (all levels in one launch through a tile table).
// Streaming stencil, HBM-bound: algorithmic bytes = level pixels read + written ("blur
// materialised" variant).
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kBW = 64, kBH = 16;          // output tile
constexpr int kRawW = kBW + 8;             // 72: 3 left + 64 + 3 right, padded to a multiple of 4
constexpr int kRawH = kBH + 6;
}

__global__ void __launch_bounds__(256)
blur_kernel(const Geom g, const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur) {
    __shared__ __align__(16) uint8_t raw[kRawH][kRawW];
    __shared__ __align__(16) uint16_t hb[kRawH][kBW];

    const int img = blockIdx.y;
    int t = blockIdx.x, l = 0;
    while (l + 1 < g.nlevels && t >= g.lv[l + 1].btile_base) ++l;
    const LevelGeom& L = g.lv[l];
    t -= L.btile_base;
    const int ty = t / L.btiles_x, tx = t - ty * L.btiles_x;
    const int x0 = tx * kBW, y0 = ty * kBH;

    // raw tile: ROI coords [x0-3, x0+69) x [y0-3, y0+19), clamped into the padded frame (clamped taps feed
    // only outputs outside the ROI, which are never stored)
    const uint8_t* src = pyr + L.off + (size_t)img * L.img_stride + (size_t)kEdge * L.pitch + kXPad;
    for (int i = threadIdx.x; i < kRawH * kRawW; i += 256) {
        const int ry = i / kRawW, rx = i - ry * kRawW;
        const int y = min(y0 - 3 + ry, L.h + kEdge - 1), x = min(x0 - 3 + rx, L.w + kEdge - 1);
        raw[ry][rx] = __ldg(src + (ptrdiff_t)y * L.pitch + x);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kRawH * kBW; i += 256) {
        const int ry = i / kBW, x = i - ry * kBW;
        const uint8_t* r = &raw[ry][x];
        hb[ry][x] = (uint16_t)(18 * (r[0] + r[6]) + 34 * (r[1] + r[5]) + 48 * (r[2] + r[4]) + 56 * r[3]);
    }
    __syncthreads();
    uint8_t* dst = blur + L.boff + (size_t)img * L.bimg_stride;
    for (int i = threadIdx.x; i < kBH * kBW / 4; i += 256) {
        const int y = i / (kBW / 4), x4 = (i - y * (kBW / 4)) * 4;
        if (y0 + y >= L.h || x0 + x4 >= L.w) continue;
        uint32_t word = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = x4 + k;
            const uint32_t v = 18u * (hb[y][x] + hb[y + 6][x]) + 34u * (hb[y + 1][x] + hb[y + 5][x]) +
                               48u * (hb[y + 2][x] + hb[y + 4][x]) + 56u * hb[y + 3][x];
            word |= ((v + 32768u) >> 16) << (8 * k);
        }
        // bpitch is a multiple of 64: the last word of a row may spill past w into row padding, never past the pitch
        *reinterpret_cast<uint32_t*>(dst + (size_t)(y0 + y) * L.bpitch + x0 + x4) = word;
    }
}

void launch_blur(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd(g.btiles_total, nimg);
    blur_kernel<<<grd, 256, 0, s>>>(g, b.pyr, b.blur);
}

}  // namespace gforb
