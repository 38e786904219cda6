// Scale pyramid: ORBextractor::ComputePyramid (reference src/ORBextractor.cc:1176-1201).
//
// Level 0 = input copied into a padded buffer; level l = cv::resize(level l-1, INTER_LINEAR) in OpenCV's
// 11-bit fixed point, cascaded; every level gets a 19-px BORDER_REFLECT_101 frame.  The frame is a pure
// function of the ROI, so each kernel writes the WHOLE padded level in one pass: a frame pixel is the
// resize evaluated at the reflected coordinate.  The source level is itself padded, so the +1 taps never
// leave valid memory and need no clamp (their weight is 0 exactly when OpenCV clamps).
//
// Memory: one thread produces 4 horizontally adjacent bytes and stores one 32-bit word; pitch is a multiple
// of 64 and the ROI starts 32 bytes into the row, so stores are aligned and coalesced.  HBM-bound:
// algorithmic bytes per image = W*H read + sum_l (w_l+38)(h_l+38) written.
#include "gforb_internal.cuh"

namespace gforb {

__device__ __forceinline__ int reflect101(int p, int len) {
    // one reflection suffices: |p| offsets are at most 19 < len for every level we accept
    p = p < 0 ? -p : p;
    return p >= len ? 2 * (len - 1) - p : p;
}

__global__ void __launch_bounds__(256)
pyr_level0_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int W, int H, int pitch, int rows,
                  size_t img_stride, int img0) {
    const int img = img0 + blockIdx.z;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    const int px4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (py >= rows || px4 >= pitch) return;
    const uint8_t* src = in + (size_t)img * W * H + (size_t)reflect101(py - kEdge, H) * W;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = px4 + k - kXPad;
        uint32_t v = 0;
        if (x >= -kEdge && x < W + kEdge) v = __ldg(src + reflect101(x, W));
        word |= v << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(out + (size_t)img * img_stride + (size_t)py * pitch + px4) = word;
}

// tabs: for x: w entries of {ofs, (a1<<16)|a0}; for y: h entries likewise (int pairs)
__global__ void __launch_bounds__(256)
pyr_resize_kernel(const uint8_t* __restrict__ src_roi, int spitch, size_t simg_stride,
                  uint8_t* __restrict__ out, int w, int h, int pitch, int rows, size_t img_stride,
                  const int2* __restrict__ tab_x, const int2* __restrict__ tab_y, int area2x, int img0) {
    const int img = img0 + blockIdx.z;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    const int px4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (py >= rows || px4 >= pitch) return;
    const uint8_t* s = src_roi + (size_t)img * simg_stride;
    const int ry = reflect101(py - kEdge, h);
    uint32_t word = 0;
    if (area2x) {
        const uint8_t* r0 = s + (size_t)(2 * ry) * spitch;
        const uint8_t* r1 = r0 + spitch;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = px4 + k - kXPad;
            uint32_t v = 0;
            if (x >= -kEdge && x < w + kEdge) {
                const int rx = 2 * reflect101(x, w);
                v = (uint32_t)(__ldg(r0 + rx) + __ldg(r0 + rx + 1) + __ldg(r1 + rx) + __ldg(r1 + rx + 1) + 2) >> 2;
            }
            word |= v << (8 * k);
        }
    } else {
        const int2 ty = __ldg(tab_y + ry);
        const int b0 = (short)(ty.y & 0xFFFF), b1 = ty.y >> 16;
        const uint8_t* r0 = s + (size_t)ty.x * spitch;
        const uint8_t* r1 = r0 + spitch;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = px4 + k - kXPad;
            uint32_t v = 0;
            if (x >= -kEdge && x < w + kEdge) {
                const int2 tx = __ldg(tab_x + reflect101(x, w));
                const int a0 = (short)(tx.y & 0xFFFF), a1 = tx.y >> 16;
                const int S0 = __ldg(r0 + tx.x) * a0 + __ldg(r0 + tx.x + 1) * a1;
                const int S1 = __ldg(r1 + tx.x) * a0 + __ldg(r1 + tx.x + 1) * a1;
                int r = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2;
                v = (uint32_t)min(max(r, 0), 255);
            }
            word |= v << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(out + (size_t)img * img_stride + (size_t)py * pitch + px4) = word;
}

void launch_pyramid(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    const dim3 blk(32, 8);
    {
        const LevelGeom& L = g.lv[0];
        dim3 grd((L.pitch / 4 + blk.x - 1) / blk.x, (L.rows + blk.y - 1) / blk.y, nimg);
        pyr_level0_kernel<<<grd, blk, 0, s>>>(b.in, b.pyr + L.off, g.W, g.H, L.pitch, L.rows, L.img_stride, g.img0);
    }
    for (int l = 1; l < g.nlevels; ++l) {
        const LevelGeom& P = g.lv[l - 1];
        const LevelGeom& L = g.lv[l];
        const uint8_t* src_roi = b.pyr + P.off + (size_t)kEdge * P.pitch + kXPad;
        dim3 grd((L.pitch / 4 + blk.x - 1) / blk.x, (L.rows + blk.y - 1) / blk.y, nimg);
        pyr_resize_kernel<<<grd, blk, 0, s>>>(src_roi, P.pitch, P.img_stride, b.pyr + L.off, L.w, L.h, L.pitch,
                                              L.rows, L.img_stride,
                                              reinterpret_cast<const int2*>(b.tabs + L.tab_x),
                                              reinterpret_cast<const int2*>(b.tabs + L.tab_y), L.area2x, g.img0);
    }
}

}  // namespace gforb
