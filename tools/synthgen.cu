// tools/synthgen.cu -- seeded synthetic rectified stereo frames, generated ON THE DEVICE (bench input, not product code).
//
// SURVEY.md 8(d) asks for distinct frames per seed (config 4: 65,536 of them): smooth 3-octave value-noise background,
// 400 discs / rotated / axis-aligned rectangles of uniform grey (FAST corners), pixel noise of sigma 2; the right image
// is the left one warped horizontally by a smooth disparity field of 2..60 px (bilinear, re-quantised) with its own
// noise.  tests/synth.py is that recipe in numpy (0.16 s per frame: fine for tests, 3 hours for config 4); this file
// is the same recipe in INTEGER arithmetic with a counter-based hash as the random source (seed = 1234 + frame), so the
// device kernels and the host loop below produce identical bytes -- the CPU arm of bench.py times the very frames the
// GPU arm does.  Built by tools/Makefile into tools/_build/libsynthgen.so; bound by tools/synthgen.py.
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <thread>
#include <vector>

#define HD __host__ __device__ __forceinline__

namespace {

constexpr int kShapes = 400;
constexpr int kTileW = 64, kTileH = 16;

struct Shape { int cx, cy, hu, hv, kind, c, s, g, reach; };

__constant__ int d_cos64[64] = {1024, 1019, 1004, 980, 946, 903, 851, 792, 724, 650, 569, 483, 392, 297, 200, 100, 0, -100, -200, -297, -392, -483, -569, -650, -724, -792, -851, -903, -946, -980, -1004, -1019, -1024, -1019, -1004, -980, -946, -903, -851, -792, -724, -650, -569, -483, -392, -297, -200, -100, 0, 100, 200, 297, 392, 483, 569, 650, 724, 792, 851, 903, 946, 980, 1004, 1019};
const int h_cos64[64] = {1024, 1019, 1004, 980, 946, 903, 851, 792, 724, 650, 569, 483, 392, 297, 200, 100, 0, -100, -200, -297, -392, -483, -569, -650, -724, -792, -851, -903, -946, -980, -1004, -1019, -1024, -1019, -1004, -980, -946, -903, -851, -792, -724, -650, -569, -483, -392, -297, -200, -100, 0, 100, 200, 297, 392, 483, 569, 650, 724, 792, 851, 903, 946, 980, 1004, 1019};

HD uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
HD uint32_t rnd(uint64_t seed, uint32_t stream, uint32_t idx) {
    return (uint32_t)(mix64(mix64(seed) ^ (((uint64_t)stream << 32) | idx)) >> 32);
}

HD Shape make_shape(uint64_t seed, int i, int w, int h, const int* cos64) {
    Shape sh;
    sh.cx = (int)(rnd(seed, 100, i) % (uint32_t)w);
    sh.cy = (int)(rnd(seed, 101, i) % (uint32_t)h);
    const int size = 4 + (int)(rnd(seed, 102, i) % 36u);
    sh.g = (int)(rnd(seed, 103, i) & 255u);
    const int k = (int)(rnd(seed, 104, i) % 100u);
    const int a = (int)(rnd(seed, 105, i) & 63u);
    const int frac = (int)(rnd(seed, 106, i) & 0xffffu);
    sh.kind = k < 30 ? 0 : (k < 65 ? 1 : 2);
    sh.hu = size / 2;
    sh.hv = sh.kind == 0 ? size / 2 : (sh.kind == 1 ? (size / 3 > 2 ? size / 3 : 2) : ((size * frac) >> 16 > 2 ? (size * frac) >> 16 : 2));
    sh.c = cos64[a];
    sh.s = cos64[(a + 48) & 63];          // sin(t) = cos(t - pi/2)
    sh.reach = size;                      // every form lies inside |dx|, |dy| <= size
    return sh;
}

HD bool covers(const Shape& sh, int x, int y) {
    const int dx = x - sh.cx, dy = y - sh.cy;
    if (sh.kind == 0) return dx * dx + dy * dy <= sh.hu * sh.hu;
    if (sh.kind == 2) return (dx < 0 ? -dx : dx) <= sh.hu && (dy < 0 ? -dy : dy) <= sh.hv;
    const int u = dx * sh.c + dy * sh.s, v = dy * sh.c - dx * sh.s;
    return (u < 0 ? -u : u) <= sh.hu * 1024 && (v < 0 ? -v : v) <= sh.hv * 1024;
}

// one octave of value noise at pixel (x, y): 16-bit result
HD uint32_t lattice_noise(uint64_t seed, uint32_t stream, int n, int x, int y, int w, int h) {
    const uint32_t px = (uint32_t)(((uint64_t)x * (uint32_t)n << 16) / (uint32_t)w);
    const uint32_t py = (uint32_t)(((uint64_t)y * (uint32_t)n << 16) / (uint32_t)h);
    const uint32_t gx = px >> 16, gy = py >> 16, fx = px & 0xffffu, fy = py & 0xffffu;
    const uint64_t a = rnd(seed, stream, gy * 64 + gx) >> 16, b = rnd(seed, stream, gy * 64 + gx + 1) >> 16;
    const uint64_t c = rnd(seed, stream, (gy + 1) * 64 + gx) >> 16, d = rnd(seed, stream, (gy + 1) * 64 + gx + 1) >> 16;
    const uint64_t top = a * (65536 - fx) + b * fx, bot = c * (65536 - fx) + d * fx;      // < 2^32
    return (uint32_t)((top * (65536 - fy) + bot * fy) >> 32);
}

HD int background(uint64_t seed, int x, int y, int w, int h) {
    uint32_t acc = 0;
    for (int o = 0; o < 3; ++o) acc += (60u >> o) * lattice_noise(seed, 10 + o, 4 << o, x, y, w, h);
    return 40 + (int)(acc >> 16);
}

// disparity in 1/256 px, 2..60 px
HD int disparity_fp8(uint64_t seed, int x, int y, int w, int h) {
    const uint32_t v = (2 * lattice_noise(seed, 20, 4, x, y, w, h) + lattice_noise(seed, 21, 8, x, y, w, h)) / 3;
    return 2 * 256 + (int)(((uint64_t)(58 * 256) * v) >> 16);
}

HD int pixel_noise(uint64_t seed, uint32_t stream, uint32_t idx) {      // binomial(16, 1/2) - 8: sigma 2
    const uint32_t r = rnd(seed, stream, idx) & 0xffffu;
#ifdef __CUDA_ARCH__
    return __popc(r) - 8;
#else
    return __builtin_popcount(r) - 8;
#endif
}

HD uint8_t clip8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

HD void finish_row(uint64_t seed, int y, int w, int h, const uint8_t* base, uint8_t* left, uint8_t* right, int x) {
    const int xs = x * 256 + disparity_fp8(seed, x, y, w, h);
    int x0 = xs >> 8;
    const int f = xs & 255;
    if (x0 > w - 1) x0 = w - 1;
    const int x1 = x0 + 1 > w - 1 ? w - 1 : x0 + 1;
    const int r = (base[x0] * (256 - f) + base[x1] * f + 128) >> 8;
    const uint32_t idx = (uint32_t)y * (uint32_t)w + (uint32_t)x;
    right[x] = clip8(r + pixel_noise(seed, 51, idx));
    left[x] = clip8((int)base[x] + pixel_noise(seed, 50, idx));
}

// ---- device ----
__global__ void __launch_bounds__(256) synth_base_kernel(uint64_t seed0, int w, int h, uint8_t* left, size_t stride) {
    __shared__ Shape sh[kShapes];
    __shared__ uint16_t list[kShapes];
    __shared__ int nlist;
    const uint64_t seed = seed0 + blockIdx.z;
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
    for (int i = threadIdx.x; i < kShapes; i += blockDim.x) sh[i] = make_shape(seed, i, w, h, d_cos64);
    __syncthreads();
    if (threadIdx.x < 32) {
        int n = 0;
        for (int k = 0; k < kShapes; k += 32) {
            const int i = k + threadIdx.x;
            bool hit = false;
            if (i < kShapes) {
                const Shape& s = sh[i];
                hit = s.cx + s.reach >= tx0 && s.cx - s.reach < tx0 + kTileW && s.cy + s.reach >= ty0 && s.cy - s.reach < ty0 + kTileH;
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) list[n + __popc(m & ((1u << threadIdx.x) - 1))] = (uint16_t)i;
            n += __popc(m);
        }
        if (threadIdx.x == 0) nlist = n;
    }
    __syncthreads();
    const int n = nlist;
    const int x = tx0 + (threadIdx.x & 63);
    uint8_t* img = left + stride * blockIdx.z;
    for (int r = threadIdx.x >> 6; r < kTileH; r += 4) {
        const int y = ty0 + r;
        if (x >= w || y >= h) continue;
        int v = -1;
        for (int k = n - 1; k >= 0; --k) {              // painter's order: the last shape drawn wins
            const Shape& s = sh[list[k]];
            if (covers(s, x, y)) { v = s.g; break; }
        }
        if (v < 0) v = background(seed, x, y, w, h);
        img[(size_t)y * w + x] = clip8(v);
    }
}

__global__ void __launch_bounds__(256) synth_finish_kernel(uint64_t seed0, int w, int h, uint8_t* left, size_t stride_l,
                                                           uint8_t* right, size_t stride_r) {
    extern __shared__ uint8_t row[];
    const uint64_t seed = seed0 + blockIdx.y;
    const int y = blockIdx.x;
    uint8_t* l = left + stride_l * blockIdx.y + (size_t)y * w;
    uint8_t* r = right + stride_r * blockIdx.y + (size_t)y * w;
    for (int x = threadIdx.x; x < w; x += blockDim.x) row[x] = l[x];
    __syncthreads();
    for (int x = threadIdx.x; x < w; x += blockDim.x) finish_row(seed, y, w, h, row, l, r, x);
}

// ---- host ----
void host_frame(uint64_t seed, int w, int h, uint8_t* left, uint8_t* right) {
    std::vector<Shape> sh(kShapes);
    for (int i = 0; i < kShapes; ++i) sh[i] = make_shape(seed, i, w, h, h_cos64);
    std::vector<uint8_t> base((size_t)w * h);
    std::vector<int> list;
    for (int ty0 = 0; ty0 < h; ty0 += kTileH)
        for (int tx0 = 0; tx0 < w; tx0 += kTileW) {
            list.clear();
            for (int i = 0; i < kShapes; ++i) {
                const Shape& s = sh[i];
                if (s.cx + s.reach >= tx0 && s.cx - s.reach < tx0 + kTileW && s.cy + s.reach >= ty0 && s.cy - s.reach < ty0 + kTileH) list.push_back(i);
            }
            for (int y = ty0; y < ty0 + kTileH && y < h; ++y)
                for (int x = tx0; x < tx0 + kTileW && x < w; ++x) {
                    int v = -1;
                    for (int k = (int)list.size() - 1; k >= 0; --k)
                        if (covers(sh[list[k]], x, y)) { v = sh[list[k]].g; break; }
                    if (v < 0) v = background(seed, x, y, w, h);
                    base[(size_t)y * w + x] = clip8(v);
                }
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) finish_row(seed, y, w, h, &base[(size_t)y * w], left + (size_t)y * w, right + (size_t)y * w, x);
}

}  // namespace

extern "C" {

// frames seed0 .. seed0+n-1 into device memory: left image f at left + f*stride_l, right at right + f*stride_r
// (tight rows of w bytes).  Asynchronous on `stream`.  Returns a cudaError_t.
int synthgen_device(uint64_t seed0, int n, int w, int h, uint8_t* left, size_t stride_l, uint8_t* right, size_t stride_r,
                    void* stream) {
    if (n < 1 || w < 1 || h < 1 || w > 8192) return (int)cudaErrorInvalidValue;
    cudaStream_t s = (cudaStream_t)stream;
    for (int f0 = 0; f0 < n; f0 += 32768) {
        const int m = n - f0 < 32768 ? n - f0 : 32768;
        dim3 g1((w + kTileW - 1) / kTileW, (h + kTileH - 1) / kTileH, m);
        synth_base_kernel<<<g1, 256, 0, s>>>(1234 + seed0 + f0, w, h, left + stride_l * f0, stride_l);
        dim3 g2(h, m);
        synth_finish_kernel<<<g2, 256, w, s>>>(1234 + seed0 + f0, w, h, left + stride_l * f0, stride_l, right + stride_r * f0, stride_r);
    }
    return (int)cudaGetLastError();
}

// the same bytes on the host cores
int synthgen_host(uint64_t seed0, int n, int w, int h, uint8_t* left, size_t stride_l, uint8_t* right, size_t stride_r,
                  int nthreads) {
    if (n < 1 || w < 1 || h < 1) return 1;
    std::atomic<int> next{0};
    auto work = [&] {
        for (;;) {
            const int f = next.fetch_add(1);
            if (f >= n) break;
            host_frame(1234 + seed0 + f, w, h, left + stride_l * f, right + stride_r * f);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < (nthreads > 1 ? nthreads : 1); ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    return 0;
}

}  // extern "C"
