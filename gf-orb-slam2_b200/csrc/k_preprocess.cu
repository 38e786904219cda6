// Input pre-processing in front of ORBextractor::operator(), one fused pass per image:
//   [16-bit -> high byte]  ->  stereo rectification  ->  colour -> gray
// Reference call sites: cv::imread(IMREAD_GRAYSCALE) of 16-bit PNGs (test/test_Stereo.cpp:100-106);
// cv::remap(im, im_rect, M1, M2, cv::INTER_LINEAR) with CV_32F maps and the default BORDER_CONSTANT 0
// (Examples/ROS/GF_ORB_SLAM2/src/ros_stereo.cc:290-295, test/test_Stereo.cpp:197-210);
// cv::cvtColor(mImGray, mImGray, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) (src/Tracking.cc:366-382).
// OpenCV arithmetic (restated from its published algorithm, pinned against cv2 in tests/test_oracle_cv2.py):
//   remap 8U bilinear = 5 fractional bits per axis: s = cvRound(map * 32); tap weights (32-fx)(32-fy) ... in Q10 after
//   dropping the common factor 32 of OpenCV's Q15 table; round by +512 >> 10; taps outside the source read 0;
//   gray = (R * 9798 + G * 19235 + B * 3735 + 16384) >> 15.
// The remap runs per channel BEFORE the gray conversion (the reference rectifies in the caller, converts in Tracking).
//
// Mapping: gather.  Thread = 4 consecutive output pixels (one 32-bit store, two 128-bit map loads); the float maps are
// shared by every image of the batch (blockIdx.y) and stay in L2 (2.9 MB at 752x480).  Algorithmic bytes per image =
// source bytes + w*h written (+ 8 B/px of maps once per batch).
#include "gforb_internal.cuh"

namespace gforb {

namespace {
// 32-bit offsets from the image base (images are < 4 GB): one IMAD per tap instead of 64-bit address arithmetic
template <int kFmt>
__device__ __forceinline__ int load_tap(const uint8_t* __restrict__ src, uint32_t pitch, int x, int y, int ch) {
    const uint32_t row = (uint32_t)y * pitch;
    if (kFmt == 0) return __ldg(src + (row + (uint32_t)x));
    if (kFmt == 1) return __ldg(reinterpret_cast<const uint16_t*>(src + (row + 2u * (uint32_t)x))) >> 8;
    if (kFmt == 2 || kFmt == 3) return __ldg(src + (row + 3u * (uint32_t)x + (uint32_t)ch));
    return __ldg(src + (row + 4u * (uint32_t)x + (uint32_t)ch));
}

template <int kFmt>
__device__ __forceinline__ int to_gray(const int (&v)[3]) {
    if (kFmt <= 1) return v[0];
    const bool bgr = kFmt == 3 || kFmt == 5;
    const int r = bgr ? v[2] : v[0], b = bgr ? v[0] : v[2];
    return (r * 9798 + v[1] * 19235 + b * 3735 + 16384) >> 15;
}

// One output pixel.  Measured on B200 (512 images 752x480, tools/bench_preprocess.py): this per-pixel form with a branch
// around the border case runs 0.47 ms; a branch-free form that issues all 16 gathers of a thread back to back 0.55 ms;
// eight images per thread sharing one map read 1.0 ms (215 registers).  ncu: the kernel is L1TEX-bound on the byte
// gathers (4 LDG.U8 wavefronts per pixel), not on DRAM, L2 or issue slots -- wider aligned loads are the next step.
template <int kFmt>
__device__ __forceinline__ int remap_pixel(const uint8_t* __restrict__ src, int sw, int sh, uint32_t pitch, float mx, float my) {
    constexpr int nch = kFmt <= 1 ? 1 : 3;
    const int sx = __float2int_rn(__fmul_rn(mx, 32.0f)), sy = __float2int_rn(__fmul_rn(my, 32.0f));     // cvRound
    const int ix = min(max(sx >> 5, -32768), 32767), iy = min(max(sy >> 5, -32768), 32767);            // saturate_cast<short>
    const int fx = sx & 31, fy = sy & 31;
    const int w00 = (32 - fx) * (32 - fy), w01 = fx * (32 - fy), w10 = (32 - fx) * fy, w11 = fx * fy;
    int v[3];
    if ((unsigned)ix < (unsigned)(sw - 1) && (unsigned)iy < (unsigned)(sh - 1)) {                       // all four taps inside
#pragma unroll
        for (int c = 0; c < nch; ++c)
            v[c] = (w00 * load_tap<kFmt>(src, pitch, ix, iy, c) + w01 * load_tap<kFmt>(src, pitch, ix + 1, iy, c) +
                    w10 * load_tap<kFmt>(src, pitch, ix, iy + 1, c) + w11 * load_tap<kFmt>(src, pitch, ix + 1, iy + 1, c) + 512) >> 10;
    } else {                                                                                            // outside taps read the constant 0
        const bool x0 = ix >= 0 && ix < sw, x1 = ix + 1 >= 0 && ix + 1 < sw, y0 = iy >= 0 && iy < sh, y1 = iy + 1 >= 0 && iy + 1 < sh;
#pragma unroll
        for (int c = 0; c < nch; ++c) {
            int a = 0;
            if (x0 && y0) a += w00 * load_tap<kFmt>(src, pitch, ix, iy, c);
            if (x1 && y0) a += w01 * load_tap<kFmt>(src, pitch, ix + 1, iy, c);
            if (x0 && y1) a += w10 * load_tap<kFmt>(src, pitch, ix, iy + 1, c);
            if (x1 && y1) a += w11 * load_tap<kFmt>(src, pitch, ix + 1, iy + 1, c);
            v[c] = (a + 512) >> 10;
        }
    }
    return to_gray<kFmt>(v);
}
}  // namespace

struct PreprocArgs {
    const uint8_t* src; size_t spitch, sstride; int sw, sh;
    const float* mapx; const float* mapy;          // dw x dh, tight; nullptr = no remap (dw x dh == sw x sh)
    uint8_t* dst; size_t dpitch, dstride; int dw, dh;
};

template <int kFmt>
__global__ void __launch_bounds__(256)
preprocess_kernel(const PreprocArgs a) {
    const int groups = (a.dw + 3) >> 2;
    const int gid = blockIdx.x * 256 + threadIdx.x;
    if (gid >= groups * a.dh) return;
    const int y = gid / groups, x0 = (gid - y * groups) << 2;
    const uint8_t* src = a.src + (size_t)blockIdx.y * a.sstride;
    uint8_t* drow = a.dst + (size_t)blockIdx.y * a.dstride + (size_t)y * a.dpitch;
    const int n = min(4, a.dw - x0);
    int g[4] = {0, 0, 0, 0};
    if (a.mapx) {
        const size_t mo = (size_t)y * a.dw + x0;
        float mx[4], my[4];
        if (n == 4 && ((mo & 3) == 0)) {
            const float4 fx = __ldg(reinterpret_cast<const float4*>(a.mapx + mo)), fy = __ldg(reinterpret_cast<const float4*>(a.mapy + mo));
            mx[0] = fx.x; mx[1] = fx.y; mx[2] = fx.z; mx[3] = fx.w; my[0] = fy.x; my[1] = fy.y; my[2] = fy.z; my[3] = fy.w;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) { mx[i] = i < n ? __ldg(a.mapx + mo + i) : 0.f; my[i] = i < n ? __ldg(a.mapy + mo + i) : 0.f; }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < n) g[i] = remap_pixel<kFmt>(src, a.sw, a.sh, (uint32_t)a.spitch, mx[i], my[i]);
    } else {
        constexpr int nch = kFmt <= 1 ? 1 : 3;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < n) {
                int v[3];
#pragma unroll
                for (int c = 0; c < nch; ++c) v[c] = load_tap<kFmt>(src, (uint32_t)a.spitch, x0 + i, y, c);
                g[i] = to_gray<kFmt>(v);
            }
    }
    if (n == 4 && ((reinterpret_cast<uintptr_t>(drow + x0) & 3) == 0)) {
        *reinterpret_cast<uint32_t*>(drow + x0) = (uint32_t)g[0] | ((uint32_t)g[1] << 8) | ((uint32_t)g[2] << 16) | ((uint32_t)g[3] << 24);
    } else {
        for (int i = 0; i < n; ++i) drow[x0 + i] = (uint8_t)g[i];
    }
}

int preprocess_bytes_per_pixel(int fmt) {
    static const int bpp[6] = {1, 2, 3, 3, 4, 4};
    return fmt >= 0 && fmt < 6 ? bpp[fmt] : 0;
}

void launch_preprocess(const uint8_t* src, int sw, int sh, size_t spitch, size_t sstride, int fmt, const float* mapx,
                       const float* mapy, uint8_t* dst, int dw, int dh, size_t dpitch, size_t dstride, int nimg, cudaStream_t s) {
    if (nimg <= 0 || dw <= 0 || dh <= 0) return;
    PreprocArgs a{src, spitch, sstride, sw, sh, mapx, mapy, dst, dpitch, dstride, dw, dh};
    const int groups = ((dw + 3) >> 2) * dh;
    dim3 grid((groups + 255) / 256, nimg);
    switch (fmt) {
        case 0: preprocess_kernel<0><<<grid, 256, 0, s>>>(a); break;
        case 1: preprocess_kernel<1><<<grid, 256, 0, s>>>(a); break;
        case 2: preprocess_kernel<2><<<grid, 256, 0, s>>>(a); break;
        case 3: preprocess_kernel<3><<<grid, 256, 0, s>>>(a); break;
        case 4: preprocess_kernel<4><<<grid, 256, 0, s>>>(a); break;
        default: preprocess_kernel<5><<<grid, 256, 0, s>>>(a); break;
    }
}

}  // namespace gforb
