// Orientation + descriptor + final keypoint assembly, one warp per kept keypoint:
//   IC_Angle / computeOrientation        (reference src/ORBextractor.cc:76-103, 472-479)  on the raw level
//   computeOrbDescriptor                 (src/ORBextractor.cc:107-146)                    on the blurred level
//   level-major concatenation, pt *= scale, size/octave fix-ups (src/ORBextractor.cc:839-849, 1141-1171)
//
// Moments are exact int32 sums (order-free), reduced with warp shuffles.  The angle is cv::fastAtan2's
// polynomial evaluated with explicitly un-fused float ops (__fmul_rn/__fadd_rn) so it is bit-identical to
// the CPU sequence.  cos/sin are taken in double and rounded to float (matches correctly-rounded cosf/sinf
// except for astronomically rare double-rounding cases; DESIGN.md "descriptor parity").  Rotated pattern
// coordinates use un-fused float ops and round-half-even (__float2int_rn == cvRound).
// Lane i computes descriptor byte i (16 gathers, 8 compares); the 32 bytes leave as one coalesced store.
// Measured alternative (round 1, B200): a warp owning 8-32 keys, doing the per-key scalar work (level lookup, fastAtan2,
// double cos/sin: ~275 of the ~960 warp-instructions per key here) once with one key per lane and only the two pixel
// passes warp-per-key, is bit-identical but SLOWER -- 0.70-0.95 ms against 0.61 ms per 512 images -- because the
// per-warp chain of dependent gathers leaves too few loads in flight at 64-125 registers; warp-per-key wins on
// memory-level parallelism.
// Also measured: the pattern as float in shared memory plus one 32-bit offset per gather -- 32 fewer I2F and ~50 fewer
// address instructions per key, yet 0.59 ms against 0.54 ms (kept: int8 pattern, 64-bit IMAD.WIDE addressing).
// Also measured (round 2): the 31 x 31 box staged in shared memory first (lane = row, three aligned LDG.128 per row, then
// lane = column reading bytes at compile-time offsets: 130 fewer instructions per key, bit-identical) -- 0.655 ms against
// 0.538 ms: the load -> store -> barrier -> load chain serialises what the 31 independent global loads overlap; CTAs
// of 4 / 16 warps instead of 8: 0.68 / 0.65 ms.
// Also measured (round 2): the table read straight into registers (two LDG.128 per lane, no per-CTA shared copy + barrier):
// 0.541 -> 0.531 ms; then as a float table, one LDG.128 per comparison (no byte extraction / I2F): 0.471 -> 0.447 ms, kept; the moments' row addresses as col +- v * pitch (one IMAD.WIDE each instead of two IADD3 of a
// running pointer): 0.577 ms, rejected -- the 30 address multiplies sit in front of the loads they feed.
// LSU/gather-bound (SURVEY.md 8d K4+K5): algorithmic bytes = 961 (31x31 raw) + 1369 (37x37 blurred) + 60 out per key.
#include <cfloat>

#include "gforb_internal.cuh"

namespace gforb {

namespace {
// The pattern as floats, laid out [comparison j of the byte][lane = byte][x0, y0, x1, y1]: one coalesced LDG.128 per
// comparison brings both points of lane i's j-th test, already converted (32 I2F + byte extractions per key less).
struct PatternF { float v[1024]; };
constexpr PatternF make_pattern_f() {
    constexpr int8_t src[1024] = {
#include "orb_pattern_31.inc"
    };
    PatternF p{};
    for (int lane = 0; lane < 32; ++lane)
        for (int j = 0; j < 8; ++j)
            for (int e = 0; e < 4; ++e) p.v[(j * 32 + lane) * 4 + e] = (float)src[32 * lane + 4 * j + e];
    return p;
}
__device__ __align__(16) const PatternF d_pattern_f = make_pattern_f();
// circular patch half-widths for HALF_PATCH_SIZE = 15 (src/ORBextractor.cc:453-468): {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10,
// 9, 8, 6, 3}, held as the nibbles of 0x3689ABCDDEEEFFFF in the kernel

#ifndef GFORB_DESC_WARPS
#define GFORB_DESC_WARPS 4      // measured with the staged box: 8 warps 0.489 ms, 4 warps 0.471, 2 warps 0.472
#endif
constexpr int kWarps = GFORB_DESC_WARPS;
#ifndef GFORB_DESC_MINB
#define GFORB_DESC_MINB 10     // CTAs per SM the register allocation must allow: 40 warps = 48 registers (32 warps: 0.538 ms, 48: 0.507)
#endif

// cv::fastAtan2 (OpenCV mathfuncs_core atan_f32), degrees
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    constexpr float sc = (float)(180.0 / 3.14159265358979323846);
    constexpr float p1 = 0.9997878412794807f * sc, p3 = -0.3258083974640975f * sc;
    constexpr float p5 = 0.1555786518463281f * sc, p7 = -0.04432655554792128f * sc;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}
}  // namespace

__global__ void __launch_bounds__(kWarps * 32, GFORB_DESC_MINB)
describe_kernel(const Geom g, const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur,
                const int32_t* __restrict__ lvl_cnt, const uint32_t* __restrict__ lvl_pts,
                gforb_keypoint* __restrict__ kps, uint8_t* __restrict__ desc, int32_t* __restrict__ nkp) {

    const int img = g.img0 + blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (slot >= g.kstage) return;

    // level of this staging slot and the keypoint counts, one level per lane (no serial search)
    const bool lv_ok = lane < g.nlevels;
    const int l = __popc(__ballot_sync(0xffffffffu, lv_ok && slot >= g.lv[lv_ok ? lane : 0].kbase)) - 1;
    const int c_l = lv_ok ? lvl_cnt[(size_t)img * g.nlevels + lane] : 0;
    const int total = __reduce_add_sync(0xffffffffu, c_l);
    const int before = __reduce_add_sync(0xffffffffu, lane < l ? c_l : 0);
    const int cnt_l = __shfl_sync(0xffffffffu, c_l, l);
    const LevelGeom& L = g.lv[l];
    const int k = slot - L.kbase;
    if (slot == 0 && lane == 0) nkp[img] = total;
    if (k >= cnt_l) return;

    const uint32_t pk = lvl_pts[(size_t)img * g.kstage + slot];
    const int x = cand_x(pk) + kMinBorder, y = cand_y(pk) + kMinBorder;   // pt += minBorder (:845-846)

    // The 37 x 37 blurred box the rotated pattern can reach (|r|, |c| <= 18), staged in shared memory with coalesced aligned
    // word loads (10 words per row from the 4-byte boundary below x - 18; lanes 0..29 take three rows per step), issued
    // before the moments so that they fly together.  The 16 gathers per lane then hit shared memory instead of costing
    // ~25 L1 wavefronts each (ncu before: L1 data pipe 77 % busy, the kernel's first limiter): 0.531 -> 0.471 ms with 4-warp CTAs at 40 warps per SM.
    __shared__ uint32_t s_box[kWarps][37 * 10];
    uint32_t* const sbx = s_box[threadIdx.x >> 5];
    const uint8_t* const bx0 = blur + L.boff + (size_t)img * L.bimg_stride + (size_t)(y - 18) * L.bpitch + (x - 18);
    const int bph = (int)((uintptr_t)bx0 & 3);
    {
        const int br = lane >= 20 ? 2 : (lane >= 10 ? 1 : 0), bw = lane - 10 * br;       // lane -> (row of the step, word)
        const int bpw = L.bpitch >> 2;
        const uint32_t* gsrc = reinterpret_cast<const uint32_t*>(bx0 - bph) + (size_t)br * bpw + bw;
        uint32_t* sdst = sbx + br * 10 + bw;
#pragma unroll
        for (int it = 0; it < 13; ++it) {
            if (lane < 30 && 3 * it + br < 37) *sdst = __ldg(gsrc);
            gsrc += 3 * bpw; sdst += 30;
        }
    }
    // ---- intensity centroid over the 749-px disc ----
    // The 31 x 31 raw box is staged like the blurred one (9 aligned words per row from the 4-byte boundary below x - 15; lanes
    // 0..26 take three rows per step, one running pointer per lane).  Then lane = column u; the disc is symmetric (umax[|u|] is
    // also the last row of column u), so the column's rows are |v| <= umax[|u|]: 31 predicated LDS.U8 at compile-time row
    // offsets.  m10 = u * column sum, m01 = sum_v v * (I(+v) - I(-v)): exact int32, order-free.
    // (The direct form -- 31 predicated LDG.U8 per lane -- cost three to four address instructions per load: the compiler
    // re-derived every row address from the kernel argument, LDC.64 + 64-bit add, to stay inside 48 registers.)
    __shared__ uint32_t s_raw[kWarps][31 * 9];
    uint32_t* const srw = s_raw[threadIdx.x >> 5];
    const int pitch = L.pitch;
    const uint8_t* const rx0 = pyr + L.off + (size_t)img * L.img_stride + (size_t)(kEdge + y - kHalfPatch) * pitch + kXPad + (x - kHalfPatch);
    const int rph = (int)((uintptr_t)rx0 & 3);
    {
        const int rr = lane >= 18 ? 2 : (lane >= 9 ? 1 : 0), rw = lane - 9 * rr;       // lane -> (row of the step, word)
        const int ppw = pitch >> 2;
        const uint32_t* gsrc = reinterpret_cast<const uint32_t*>(rx0 - rph) + (size_t)rr * ppw + rw;
        uint32_t* sdst = srw + rr * 9 + rw;
#pragma unroll
        for (int it = 0; it < 11; ++it) {
            if (lane < 27 && 3 * it + rr < 31) *sdst = __ldg(gsrc);
            gsrc += 3 * ppw; sdst += 27;
        }
    }
    __syncwarp();
    const int u = lane - kHalfPatch;
    int m10 = 0, m01 = 0;
    {
        // umax[|u|] out of a 64-bit immediate (4 bits per entry) instead of a lane-indexed (divergent) __constant__ load
        const int vl = lane < 31 ? (int)((0x3689ABCDDEEEFFFFull >> (4 * abs(u))) & 15ull) : -1;
        const uint8_t* col = reinterpret_cast<const uint8_t*>(srw) + kHalfPatch * 36 + min(lane, 30) + rph;    // (row 0, column u)
        int up[kHalfPatch + 1], dn[kHalfPatch + 1];
        up[0] = 0 <= vl ? (int)col[0] : 0;
        dn[0] = 0;
#pragma unroll
        for (int v = 1; v <= kHalfPatch; ++v) {
            up[v] = v <= vl ? (int)col[-36 * v] : 0;
            dn[v] = v <= vl ? (int)col[36 * v] : 0;
        }
        int sum = up[0];
#pragma unroll
        for (int v = 1; v <= kHalfPatch; ++v) {
            sum += up[v] + dn[v];
            m01 += v * (dn[v] - up[v]);
        }
        m10 = u * sum;
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    m01 = __reduce_add_sync(0xffffffffu, m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rBRIEF on the blurred level ----
    const float factorPI = (float)(3.14159265358979323846 / 180.0);
    const float ang = __fmul_rn(angle, factorPI);
    double sd, cd;
    sincos((double)ang, &sd, &cd);                     // one shared range reduction; each result correctly rounded to float below
    const float a = (float)cd, b = (float)sd;
    // lane i computes descriptor byte i from pattern points 16 i .. 16 i + 15
    const float4* const pat = reinterpret_cast<const float4*>(d_pattern_f.v) + lane;
    __syncwarp();                                      // the staged box is complete
    uint32_t val = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        int t[2];
        const float4 pq = __ldg(pat + 32 * j);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float px = h ? pq.z : pq.x, py = h ? pq.w : pq.y;
            const int r = __float2int_rn(__fadd_rn(__fmul_rn(px, b), __fmul_rn(py, a)));
            const int c = __float2int_rn(__fsub_rn(__fmul_rn(px, a), __fmul_rn(py, b)));
            t[h] = reinterpret_cast<const uint8_t*>(sbx)[(r + 18) * 40 + (c + 18 + bph)];
        }
        val |= (uint32_t)(t[0] < t[1]) << j;
    }
    const size_t o = (size_t)img * g.K + before + k;
    desc[o * 32 + lane] = (uint8_t)val;

    if (lane == 0) {
        gforb_keypoint kp;
        kp.x = l ? __fmul_rn((float)x, L.scale) : (float)x;     // keypoint->pt *= scale (:1164-1170)
        kp.y = l ? __fmul_rn((float)y, L.scale) : (float)y;
        kp.size = L.patch_size;
        kp.angle = angle;
        kp.response = (float)cand_r(pk);
        kp.octave = l;
        kp.class_id = -1;
        kps[o] = kp;
    }
}

void launch_describe(const Geom& g, const Buffers& b, int nimg, cudaStream_t s) {
    dim3 grd((g.kstage + kWarps - 1) / kWarps, nimg);
    describe_kernel<<<grd, kWarps * 32, 0, s>>>(g, b.pyr, b.blur, b.lvl_cnt, b.lvl_pts, b.kps, b.desc, b.nkp);
}

void upload_constants() {}

}  // namespace gforb
