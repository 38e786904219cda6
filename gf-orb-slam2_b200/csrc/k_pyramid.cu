// Scale pyramid: ORBextractor::ComputePyramid (reference src/ORBextractor.cc:1176-1201).
//
// Level 0 = input copied into a padded buffer; level l = cv::resize(level l-1, INTER_LINEAR) in OpenCV's
// 11-bit fixed point, cascaded; every level gets a 19-px BORDER_REFLECT_101 frame.  The frame is a pure
// function of the ROI: a frame pixel is the resize evaluated at the reflected coordinate (generic kernels below: the WHOLE
// padded level in one pass; fast path: see there).  The source level is itself padded, so the +1 taps never
// leave valid memory and need no clamp (their weight is 0 exactly when OpenCV clamps).
//
// Memory: one thread produces 4 horizontally adjacent bytes and stores one 32-bit word; pitch is a multiple
// of 64 and the ROI starts 32 bytes into the row, so stores are aligned and coalesced.
// Algorithmic bytes per image = W*H read + sum_l (w_l+38)(h_l+38) written; measured traffic is 1.57x
// that, because level l is read back from HBM by the launch of level l + 1 (a 512-image level is 50-220 MB against the
// 126 MB L2).  The kernels are instruction-issue bound (70-84 % of the issue slots, 23-50 % of the DRAM rate), so the
// re-read costs no time.  Measured and rejected (round 2): walking the images in opposite directions on alternate levels
// so that a launch starts with what the previous one wrote last (0.551 against 0.530 ms); 64-row tiles (0.571 ms).
// Late in round 2: two vertically adjacent output tiles per CTA, both source boxes requested up front (second TMA in flight
// during the first tile, per-thread set-up paid once): 0.475 against 0.464 ms (64 registers instead of 56).
#include "gforb_internal.cuh"
#include "tma.cuh"

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


// ------------------------------------------------------------------------------------------------
// Fast path (v3).  A thread owns 4 horizontally adjacent ROI pixels (one aligned 32-bit word) and walks down
// kRowsPerThread rows; a warp owns 32 adjacent groups; the last warp column is shifted left to start at group ng-32
// (the overlap recomputes identical words), so an ROI edge never straddles two warps.
//
// Who reads the 19-px BORDER_REFLECT_101 frame (copyMakeBorder after resize, src/ORBextractor.cc:1189-1197)?  The next
// level's resize does not (its taps outside the ROI carry weight 0); FAST, the orientation patch and the stereo windows
// stay >= 16 px inside the ROI; the descriptor reads the blurred level, which has no frame.  Only the 7x7 blur does, and it
// reaches 3 px.  So the level kernels write the ROI plus a kBlurFrame = 3 px frame, and the other 16 px are written by
// pyr_frame_kernel only where somebody can look at them: the single-image extractor handles, whose
// gforb_pyramid_level / gforb_download_pyramid expose the padded level (mvImagePyramid).  The batch path skips them.
// The 3-px frame is not computed either: frame pixel == ROI pixel at the reflected coordinate, so
//  * lane 0 of the left-edge warp stores {-, p3, p2, p1} (one PRMT of its own word) in the word left of the ROI;
//  * the lane of the last group stores, from (previous lane's word, own word) with two PRMT, its own word completed with
//    mirrored pixels (w % 4 != 0) and the word after it;
//  * rows 1..3 and h-4..h-2 are stored a second time at their mirror rows.
// Every store is an aligned 32-bit word; interior warps run no edge code at all.
// v2 wrote all 19 px this way (5 + 5 side words per row, 2 x 19 mirror rows): 0.534 ms per 512 images, v3 0.467 ms; measured on the way
// here: no frame at all 0.402 ms, the whole frame in a separate kernel 0.86 ms (side strips = two cold 20-byte accesses
// per row), side words in the level kernels + top / bottom rows in a separate kernel 0.515 ms.
namespace {
constexpr int kRowsPerThread = 8;                       // level 0
constexpr int kResizeRows = kResizeTileRows / 4;        // resize: rows per thread (4 warps per CTA)
static_assert(kResizeRows == 8, "the row table is fetched eight rows at a time");
constexpr int kBlurFrame = 3;                           // frame pixels the blur reads (7 taps)

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {   // selector nibbles <= 7 only
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// PRMT selectors of the last group's two words over (previous lane's word, own word): byte i of that pair = pixel
// XG - 4 + i (XG = first pixel of the last group); word 0 = columns XG..XG+3, word 1 = the following four; a column inside the
// ROI takes its pixel, a frame column up to w + 2 its REFLECT_101 image, columns past that are nobody's (byte 0)
inline uint2 edge_selectors(int w) {
    const int ng = (w + 3) >> 2, q = 2 * (w - 1), XG = 4 * (ng - 1);
    uint2 sel = make_uint2(0u, 0u);
    for (int i = 0; i < 8; ++i) {
        const int c = XG + i;
        const int src = c < w ? c : q - c;
        const uint32_t idx = c < w + kBlurFrame ? (uint32_t)(src - XG + 4) : 0u;
        if (i < 4) sel.x |= idx << (4 * i);
        else sel.y |= idx << (4 * (i - 4));
    }
    return sel;
}

struct RowStore {
    int pitch, h;
    uint32_t sel_m, sel_f;   // last group: own word / following word out of (previous lane's word, own word)
    bool valid, lon, last;   // thread owns ROI pixels / is lane 0 of the left-edge warp / owns the last group
    bool ledge, redge;       // warp-uniform: warp holds the left / right ROI edge
    bool yedge;              // warp-uniform: the strip holds rows that are mirrored into the 3-px frame
    bool plain;              // warp-uniform: none of the above

    // gx = group index (may be >= ng: idle lane kept alive for the shuffle), g0 = first group of the warp
    // sels = the level's two selectors (edge_selectors, computed on the host once per launch)
    __device__ __forceinline__ void init(int pitch_, int gx, int g0, int w, int h_, int y0, int y1, uint2 sels) {
        const int ng = (w + 3) >> 2;
        pitch = pitch_; h = h_;
        sel_m = sels.x; sel_f = sels.y;
        valid = gx < ng;
        ledge = g0 == 0;
        redge = g0 + 32 >= ng;
        yedge = y0 <= kBlurFrame || y1 - 1 >= h - 1 - kBlurFrame;
        plain = !(ledge | redge | yedge);
        lon = gx == 0;
        last = gx == ng - 1;
    }
    // r = this thread's word on some padded row
    __device__ __forceinline__ void put(uint8_t* r, uint32_t m, uint32_t lw, uint32_t rw) const {
        if (valid) *reinterpret_cast<uint32_t*>(r) = m;
        if (ledge && lon) *reinterpret_cast<uint32_t*>(r - 4) = lw;
        if (redge && last) *reinterpret_cast<uint32_t*>(r + 4) = rw;
    }
    // all 32 lanes of a non-plain warp call this together (shuffle); r = this thread's word on ROI row y
    __device__ __forceinline__ void store_at(uint8_t* r, int y, uint32_t own) const {
        uint32_t m = own, lw = 0, rw = 0;
        if (ledge) lw = prmt(own, own, 0x1230u);                           // columns -3..-1 = pixels 3, 2, 1
        if (redge) {
            const uint32_t prev = __shfl_up_sync(0xffffffffu, own, 1);
            if (last) { m = prmt(prev, own, sel_m); rw = prmt(prev, own, sel_f); }
        }
        put(r, m, lw, rw);
        if (yedge) {
            if (y >= 1 && y <= kBlurFrame) put(r - (ptrdiff_t)(2 * y) * pitch, m, lw, rw);
            if (y >= h - 1 - kBlurFrame && y <= h - 2) put(r + (ptrdiff_t)(2 * (h - 1 - y)) * pitch, m, lw, rw);
        }
    }
};
}  // namespace

// level 0: input (tight, W % 4 == 0) -> ROI + 3-px frame of the padded level; thread = 4 pixels x kRowsPerThread rows
__global__ void __launch_bounds__(128)
pyr_level0_fast_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int W, int H, int pitch,
                       size_t img_stride, int img0, uint2 sels) {
    const int img = img0 + blockIdx.z;
    const int ng = W >> 2;
    const int g0 = min((int)blockIdx.x * 32, max(ng - 32, 0));
    const int gx = g0 + threadIdx.x;
    const int y0 = (blockIdx.y * 4 + threadIdx.y) * kRowsPerThread;
    if (y0 >= H) return;
    const int y1 = min(y0 + kRowsPerThread, H);
    const uint8_t* src = in + (size_t)img * W * H + 4 * min(gx, ng - 1);
    RowStore st;
    st.init(pitch, gx, g0, W, H, y0, y1, sels);
    uint8_t* p = out + (size_t)img * img_stride + (size_t)(kEdge + y0) * pitch + kXPad + 4 * gx;
    uint32_t v[kRowsPerThread];
#pragma unroll
    for (int r = 0; r < kRowsPerThread; ++r)
        v[r] = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)min(y0 + r, H - 1) * W));
#pragma unroll
    for (int r = 0; r < kRowsPerThread; ++r) {
        if (y0 + r < y1) {
            if (st.plain) *reinterpret_cast<uint32_t*>(p) = v[r];
            else st.store_at(p, y0 + r, v[r]);
        }
        p += pitch;
    }
}

// The rest of the 19-px frame of ALL levels of an image, for the handles that expose the padded levels.  Thread = one
// aligned 32-bit word of a frame: 2 x 19 top / bottom rows (words from ROI column -20 up to the one that holds column
// w + 18), and on the h ROI rows the 5 words of columns [-20, -1] plus the words from w & ~3 to the one holding column
// w + 18.  A byte takes the ROI pixel at the reflected coordinate (bytes outside [-19, w + 19) take 0; ROI bytes of a
// partial last word are copied onto themselves; the 3-px frame of the level kernels is rewritten with the same values).
struct FrameItems { int base[kMaxLevels + 1]; };
__global__ void __launch_bounds__(256)
pyr_frame_kernel(const Geom g, uint8_t* pyr, const FrameItems fi) {
    const int img = g.img0 + blockIdx.y;
    int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= fi.base[g.nlevels]) return;
    int l = 0;
    while (i >= fi.base[l + 1]) ++l;
    i -= fi.base[l];
    const LevelGeom& L = g.lv[l];
    const int w = L.w, h = L.h, pitch = L.pitch;
    const int nW = ((w + 38) >> 2) + 1, XR0 = w & ~3, nS = 5 + ((w + 18 - XR0) >> 2) + 1;
    int py, X;                                            // padded row; ROI column of the word's first byte
    const int ntb = 2 * kEdge * nW;
    if (i < ntb) {
        const int r = i / nW;
        X = -20 + 4 * (i - r * nW);
        py = r < kEdge ? r : h + r;
    } else {
        i -= ntb;
        const int r = i / nS, k = i - r * nS;
        py = kEdge + r;
        X = k < 5 ? -20 + 4 * k : XR0 + 4 * (k - 5);
    }
    uint8_t* lvl = pyr + L.off + (size_t)img * L.img_stride;
    const uint8_t* src = lvl + (size_t)(kEdge + reflect101(py - kEdge, h)) * pitch + kXPad;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = X + k;
        if (c >= -kEdge && c < w + kEdge) word |= (uint32_t)src[reflect101(c, w)] << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(lvl + (size_t)py * pitch + kXPad + X) = word;
}

__device__ __forceinline__ void lds3(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c) {
    asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
                 : "=r"(a), "=r"(b), "=r"(c) : "r"(addr));
}
__device__ __forceinline__ void resize_hrow(uint32_t w0, uint32_t w1, uint32_t w2, const uint4& sel,
                                            const uint32_t (&cf)[4], uint32_t (&s)[4]) {
    const uint32_t A = prmt(w0, w1, sel.y);
    const uint32_t B = prmt(prmt(w0, w1, sel.z), w2, sel.w);
    s[0] = __dp2a_lo(cf[0], A, 0u) >> 4;
    s[1] = __dp2a_hi(cf[1], A, 0u) >> 4;
    s[2] = __dp2a_lo(cf[2], B, 0u) >> 4;
    s[3] = __dp2a_hi(cf[3], B, 0u) >> 4;
}

// level l >= 1 (not the exact-2x case): cv::resize INTER_LINEAR 8U, thread = 4 pixels x kResizeRows rows,
// CTA = 128 pixels x kResizeTileRows rows.  kTma: the source rows/columns the tile needs arrive as ONE TMA box
// (cp.async.bulk.tensor.3d of the padded source level) on an mbarrier; the row table entries are fetched while it
// is in flight.  Consecutive output rows share source rows (scale 1.2: the lower tap row of y is usually the
// upper tap row of y+1), so a source row's horizontal pass is reused when the row table says so.  Weights are in
// [0, 2048], so every intermediate is non-negative: (b * s) >> 16 is one IMAD.HI with b pre-shifted by 16, and
// the saturating cast of cv::resize can never clip.
template <bool kTma>
__global__ void __launch_bounds__(128)
pyr_resize_fast_kernel(const __grid_constant__ CUtensorMap smap, const uint8_t* __restrict__ src_roi, int spitch,
                       size_t simg_stride, uint8_t* __restrict__ out, int w, int h, int pitch, size_t img_stride,
                       const uint4* __restrict__ tab_g, const int2* __restrict__ tab_y, int bw, int bh, int img0, uint2 sels) {
    extern __shared__ __align__(128) uint8_t tile[];
    __shared__ __align__(8) uint64_t bar;
    const int img = img0 + blockIdx.z;
    const int ng = (w + 3) >> 2;
    const int g0 = min((int)blockIdx.x * 32, max(ng - 32, 0));
    const int gx = g0 + threadIdx.x;
    const int ty0 = blockIdx.y * kResizeTileRows;
    const int y0 = ty0 + threadIdx.y * kResizeRows;
    int xb = 0, sy0 = 0;
    if (kTma) {
        xb = (int)__ldg(tab_g + 2 * g0).x & ~15;     // tile's first source byte, 16-byte aligned (TMA rule)
        sy0 = __ldg(tab_y + ty0).x;
        if (threadIdx.x == 0 && threadIdx.y == 0) {
            mbar_init(&bar, 1);
            mbar_expect_tx(&bar, (uint32_t)(bw * bh));
            tma_load_tile_3d(tile, &smap, kXPad + xb, kEdge + sy0, img, &bar);
        }
        __syncthreads();     // barrier initialised before anyone polls it
    }
    if (y0 >= h) return;     // whole warps only: idle lanes (gx >= ng) stay for the edge shuffles
    const int y1 = min(y0 + kResizeRows, h);
    const int gc = min(gx, ng - 1);
    const uint4 ga = __ldg(tab_g + 2 * gc), gb = __ldg(tab_g + 2 * gc + 1);
    const uint32_t cf[4] = {gb.x, gb.y, gb.z, gb.w};
    RowStore st;
    st.init(pitch, gx, g0, w, h, y0, y1, sels);
    const uint8_t* s = src_roi + (size_t)img * simg_stride + ga.x;          // !kTma: global source
    const uint32_t sa = smem_u32(tile) + (uint32_t)((int)ga.x - xb);        // kTma: shared byte address of this group
    if (kTma) mbar_wait(&bar, 0);

    // One running row pointer (the address arithmetic per row was a third of this kernel's instructions), and the
    // interior warps -- no ROI edge, no frame row, every lane live: warp-uniform -- store without any edge code.
    uint8_t* p = out + (size_t)img * img_stride + (size_t)(kEdge + y0) * pitch + kXPad + 4 * gx;
    const bool plain = st.plain;
    // The two horizontal-pass results ping-pong between sA and sB: on even rows sA is the upper tap row and sB the lower,
    // on odd rows the other way round, so a reused row (the lower row of y is the upper row of y + 1 five times out of
    // six at scale 1.2) costs no register moves -- the loop is fully unrolled and the roles are static.
    uint32_t sA[4] = {0u, 0u, 0u, 0u}, sB[4] = {0u, 0u, 0u, 0u};
    int have = -1;
    int2 ty[8];
    auto hrow = [&](int sy, uint32_t (&dst)[4]) {
        uint32_t w0, w1, w2;
        if (kTma) lds3(sa + (uint32_t)((sy - sy0) * bw), w0, w1, w2);
        else {
            const uint32_t* p0 = reinterpret_cast<const uint32_t*>(s + (size_t)sy * spitch);
            w0 = __ldg(p0); w1 = __ldg(p0 + 1); w2 = __ldg(p0 + 2);
        }
        resize_hrow(w0, w1, w2, ga, cf, dst);
    };
    auto row = [&](int r, int y, uint32_t (&top)[4], uint32_t (&bot)[4]) {
        const int sy = ty[r & 7].x;
        const uint32_t b0 = (uint32_t)ty[r & 7].y << 16, b1 = (uint32_t)ty[r & 7].y & 0xFFFF0000u;
        if (sy != have) hrow(sy, top);
        hrow(sy + 1, bot);
        have = sy + 1;
        uint32_t t[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] = __umulhi(b0, top[k]) + __umulhi(b1, bot[k]) + 2u;   // <= 1022
        // (t >> 2) of pixels k, k+1 land in bytes 0 and 2 of ((t[k+1] << 16) | t[k]) >> 2
        const uint32_t u01 = (t[1] * 65536u + t[0]) >> 2, u23 = (t[3] * 65536u + t[2]) >> 2;
        const uint32_t word = prmt(u01, u23, 0x6420);
        if (plain) *reinterpret_cast<uint32_t*>(p) = word;
        else st.store_at(p, y, word);
        p += pitch;
    };
#pragma unroll
    for (int q = 0; q < 8; ++q) ty[q] = __ldg(tab_y + min(y0 + q, h - 1));
    if (y1 - y0 == kResizeRows) {            // every strip but the last of a level: no per-row bound test
#pragma unroll
        for (int r = 0; r < kResizeRows; r += 2) {
            row(r, y0 + r, sA, sB);
            row(r + 1, y0 + r + 1, sB, sA);
        }
    } else {
#pragma unroll
        for (int r = 0; r < kResizeRows; r += 2) {
            if (y0 + r < y1) row(r, y0 + r, sA, sB);
            if (y0 + r + 1 < y1) row(r + 1, y0 + r + 1, sB, sA);
        }
    }
}

void launch_pyramid(const Geom& g, const Buffers& b, int nimg, cudaStream_t s, const CUtensorMap* rsmaps, bool full_frames) {
    const dim3 blk(32, 8);
    const dim3 fblk(32, 4);
    {
        const LevelGeom& L = g.lv[0];
        if (g.W % 4 == 0) {
            dim3 grd((g.W / 4 + 31) / 32, (g.H + 4 * kRowsPerThread - 1) / (4 * kRowsPerThread), nimg);
            pyr_level0_fast_kernel<<<grd, fblk, 0, s>>>(b.input(), b.pyr + L.off, g.W, g.H, L.pitch, L.img_stride, g.img0,
                                                        edge_selectors(g.W));
        } else {
            dim3 grd((L.pitch / 4 + blk.x - 1) / blk.x, (L.rows + blk.y - 1) / blk.y, nimg);
            pyr_level0_kernel<<<grd, blk, 0, s>>>(b.input(), b.pyr + L.off, g.W, g.H, L.pitch, L.rows, L.img_stride, g.img0);
        }
    }
    for (int l = 1; l < g.nlevels; ++l) {
        const LevelGeom& P = g.lv[l - 1];
        const LevelGeom& L = g.lv[l];
        const uint8_t* src_roi = b.pyr + P.off + (size_t)kEdge * P.pitch + kXPad;
        if (L.tab_g >= 0) {
            dim3 grd(((L.w + 3) / 4 + 31) / 32, (L.h + kResizeTileRows - 1) / kResizeTileRows, nimg);
            const uint4* tg = reinterpret_cast<const uint4*>(b.tabs + L.tab_g);
            const int2* ty = reinterpret_cast<const int2*>(b.tabs + L.tab_y);
            if (rsmaps && L.rs_bw)
                pyr_resize_fast_kernel<true><<<grd, fblk, (size_t)L.rs_bw * L.rs_bh, s>>>(
                    rsmaps[l], src_roi, P.pitch, P.img_stride, b.pyr + L.off, L.w, L.h, L.pitch, L.img_stride, tg, ty,
                    L.rs_bw, L.rs_bh, g.img0, edge_selectors(L.w));
            else
                pyr_resize_fast_kernel<false><<<grd, fblk, 0, s>>>(
                    CUtensorMap{}, src_roi, P.pitch, P.img_stride, b.pyr + L.off, L.w, L.h, L.pitch, L.img_stride, tg, ty,
                    0, 0, g.img0, edge_selectors(L.w));
            continue;
        }
        dim3 grd((L.pitch / 4 + blk.x - 1) / blk.x, (L.rows + blk.y - 1) / blk.y, nimg);
        pyr_resize_kernel<<<grd, blk, 0, s>>>(src_roi, P.pitch, P.img_stride, b.pyr + L.off, L.w, L.h, L.pitch,
                                              L.rows, L.img_stride,
                                              reinterpret_cast<const int2*>(b.tabs + L.tab_x),
                                              reinterpret_cast<const int2*>(b.tabs + L.tab_y), L.area2x, g.img0);
    }
    if (full_frames) {
        FrameItems fi;
        fi.base[0] = 0;
        for (int l = 0; l < kMaxLevels; ++l) {
            fi.base[l + 1] = fi.base[l];
            if (l >= g.nlevels) continue;
            const int w = g.lv[l].w, h = g.lv[l].h;
            const int nW = ((w + 38) >> 2) + 1, nS = 5 + ((w + 18 - (w & ~3)) >> 2) + 1;
            fi.base[l + 1] += 2 * kEdge * nW + h * nS;
        }
        dim3 grd((fi.base[g.nlevels] + 255) / 256, nimg);
        pyr_frame_kernel<<<grd, 256, 0, s>>>(g, b.pyr, fi);
    }
}

}  // namespace gforb
