// Stereo association: Frame::ComputeStereoMatches (reference src/Frame.cc:889-1078) and
// Frame::ComputeStereoMatches_Undistorted + PrepareStereoCandidates (src/Frame.cc:1167-1316,
// include/Frame.h:230-263), with ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1768-1784) as __popc.
//
// kSlices CTAs per stereo frame (left keys are dealt round-robin to the slices; the last CTA of a frame to
// finish -- ticket counter -- applies the median cut).  The reference builds a per-row table of right keys and scans it per left key;
// since "best" is the lexicographic minimum of (Hamming distance, right index) over the gated candidates
// (strict '<' while walking ascending iR), the table is replaced by a gate test of every right key
// (row band, octave +-1, uR window) and a warp-wide packed-min reduction -- order-free and deterministic.
// Right-key gate data sits in shared memory; descriptors are 2 x 128-bit loads per side.
// Refinement (classic mode): 11x11 SAD at 11 shifts on the un-blurred pyramids at the LEFT key's octave,
// exact in integers (the reference's float sums are integer-valued), parabola in un-fused float ops.
// The median cut needs only the value of the (n/2)-th smallest distance: rank selection in shared memory.
// INT/LSU-bound.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kT = 256;
#ifndef GFORB_ST_SLICES
#define GFORB_ST_SLICES 8
#endif
#ifndef GFORB_ST_MINB
#define GFORB_ST_MINB 4
#endif
constexpr int kSlices = GFORB_ST_SLICES;
struct RKey { short minr, maxr; uint32_t oi; float x; };   // oi = octave << 16 | original index

__device__ __forceinline__ int hamming_u4(const uint4 a0, const uint4 a1, const uint8_t* b) {
    const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(b));
    const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(b) + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
// (PTX vabsdiff.s32.add is emulated with 64-bit arithmetic on sm_100a, ~9 SASS instructions: the SAD uses IABS)
__device__ __forceinline__ RKey make_rkey(const gforb_keypoint& k, int i, const float* scale, int nRows) {
    const float r = __fmul_rn(2.0f, scale[k.octave]);
    float hi = ceilf(__fadd_rn(k.y, r)), lo = floorf(__fsub_rn(k.y, r));
    hi = fminf(hi, (float)(nRows - 1));   // classic reference indexes unclamped (keys sit >= 19*scale inside)
    lo = fmaxf(lo, 0.0f);
    RKey q;
    q.minr = (short)(int)lo; q.maxr = (short)(int)hi; q.oi = ((uint32_t)k.octave << 16) | (uint32_t)i; q.x = k.x;
    return q;
}
__device__ __forceinline__ int row_bin(float y, int nRows) { return min(max((int)floorf(y), 0), nRows - 1); }

// k-th smallest (0-based) of the non-negative entries of d[0..n) that are < 2^16, by two 256-bin histogram passes.
// All kT threads call it; hist = 256 ints of shared memory; s_tmp = 2 ints.
__device__ int select_kth(const int* d, int n, int k, int* hist, int* s_tmp) {
    int prefix = 0;
    for (int pass = 0; pass < 2; ++pass) {
        const int shift = pass ? 0 : 8;
        hist[threadIdx.x] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += kT) {
            const int e = d[i];
            if (e >= 0 && (pass == 0 || (e >> 8) == prefix)) atomicAdd(&hist[(e >> shift) & 255], 1);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            const int lane = threadIdx.x;
            int loc[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { loc[j] = hist[8 * lane + j]; sum += loc[j]; }
            int inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            int run = inc - sum;
            if (run <= k && k < inc) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (run <= k && k < run + loc[j]) { s_tmp[0] = 8 * lane + j; s_tmp[1] = k - run; }
                    run += loc[j];
                }
            }
        }
        __syncthreads();
        prefix = pass ? ((prefix << 8) | s_tmp[0]) : s_tmp[0];
        k = s_tmp[1];
        __syncthreads();
    }
    return prefix;
}
}  // namespace

__global__ void __launch_bounds__(kT, GFORB_ST_MINB)
stereo_kernel(const StereoView v, const gforb_stereo_params prm, int K, int band) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    // [rs: K sorted right keys][binstart: nRows + 1][s_dist: K (aliases rs once the matching is done)]
    RKey* rs = reinterpret_cast<RKey*>(smem_raw);
    int* binstart = reinterpret_cast<int*>(smem_raw + sizeof(RKey) * (size_t)K);
    int* s_dist = reinterpret_cast<int*>(smem_raw);
    __shared__ int s_hist[256];
    __shared__ int s_tmp[2];
    __shared__ int s_cnt, s_removed, s_processed, s_last;

    const int f = v.f0 + blockIdx.x, slice = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const gforb_keypoint* kL = v.kL + (size_t)f * v.k_stride;
    const gforb_keypoint* kR = v.kR + (size_t)f * v.k_stride;
    const uint8_t* dL = v.dL + (size_t)f * v.k_stride * 32;
    const uint8_t* dR = v.dR + (size_t)f * v.k_stride * 32;
    const int nL = min(v.nL[(size_t)f * v.n_stride], K), nR = min(v.nR[(size_t)f * v.n_stride], K);
    float* uright = v.uright + (size_t)f * v.out_stride;
    float* depth = v.depth + (size_t)f * v.out_stride;
    int32_t* o_bd = v.best_dist + (size_t)f * v.out_stride;
    int32_t* o_bi = v.best_idx + (size_t)f * v.out_stride;
    int32_t* o_sad = v.sad + (size_t)f * v.out_stride;
    int32_t* o_cut = v.cut_dist + (size_t)f * v.out_stride;

    const bool classic = prm.mode == GFORB_STEREO_CLASSIC;
    const int th_high = prm.th_high, thOrbDist = (prm.th_high + prm.th_low) / 2;
    const int nRows = v.nRows;

    // ---- right keys sorted by centre row (counting sort): a left key then scans only rows row-band .. row+band.
    // The reference's per-row candidate table (:906-916 / Frame.h:246-257) holds key i on rows minr..maxr with
    // |row - y| < 2*scale + 1, so the centre-row window is a superset and the exact gate test below decides.
    for (int i = threadIdx.x; i <= nRows; i += kT) binstart[i] = 0;
    if (threadIdx.x == 0) { s_cnt = 0; s_removed = 0; s_processed = 0; }
    __syncthreads();
    for (int i = threadIdx.x; i < nR; i += kT) atomicAdd(&binstart[row_bin(kR[i].y, nRows) + 1], 1);
    __syncthreads();
    {   // inclusive scan of binstart[1..nRows] (binstart[0] = 0): contiguous chunk per thread + block scan of chunk sums
        const int per = (nRows + kT - 1) / kT, b0 = 1 + threadIdx.x * per, b1 = min(b0 + per, nRows + 1);
        int sum = 0;
        for (int i = b0; i < b1; ++i) sum += binstart[i];
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_hist[warp] = inc;
        __syncthreads();
        int base = 0;
        for (int w = 0; w < warp; ++w) base += s_hist[w];
        int run = base + inc - sum;
        for (int i = b0; i < b1; ++i) { run += binstart[i]; binstart[i] = run; }
        __syncthreads();
    }
    // scatter from the back of each bin: binstart[b + 1] is decremented down to the bin's start, i.e. afterwards
    // binstart[b + 1] = first slot of bin b  ->  shift the view by one: bin b occupies [bs[b + 1], bs[b + 2])
    for (int i = threadIdx.x; i < nR; i += kT) {
        const gforb_keypoint k = kR[i];
        const int slot = atomicSub(&binstart[row_bin(k.y, nRows) + 1], 1) - 1;
        rs[slot] = make_rkey(k, i, v.scale, nRows);
    }
    __syncthreads();
    // now bin b = [binstart[b + 1], b + 1 < nRows ? binstart[b + 2] : nR)

    const float minZ = prm.b;
    for (int iL = slice + kSlices * warp; iL < nL; iL += kSlices * (kT / 32)) {
        const gforb_keypoint kp = kL[iL];
        float minD = 0.f, maxD = __fdiv_rn(prm.bf, minZ);
        const float vL = kp.y, uL = kp.x;
        const int levelL = kp.octave;
        float out_u = -1.f, out_z = -1.f;
        int out_bd = th_high, out_bi = -1, out_sad = -1, out_dist = -1;
        bool processed = false;

        bool ok = classic ? true : !(vL < 0 || vL > (float)(nRows - 1));
        const int row = (int)vL;
        ok = ok && row >= 0 && row < nRows;
        if (!classic && v.min_d != nullptr && v.max_d != nullptr) { minD = v.min_d[iL]; maxD = v.max_d[iL]; }
        const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);

        if (ok) {
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32) + 1);
            unsigned best = ((unsigned)th_high << 16) | 0xFFFFu;
            bool any_row = false;
            const bool range_ok = classic ? !(maxU < 0) : !(maxU < prm.min_x);
            const int blo = max(row - band, 0), bhi = min(row + band, nRows - 1);
            const int jend = bhi + 1 < nRows ? binstart[bhi + 2] : nR;
            for (int j = binstart[blo + 1] + lane; j < jend; j += 32) {
                const RKey q = rs[j];
                if (row < q.minr || row > q.maxr) continue;
                any_row = true;
                if (!range_ok) continue;
                const int oct = (int)(q.oi >> 16);
                if (oct < levelL - 1 || oct > levelL + 1) continue;
                if (q.x >= minU && q.x <= maxU) {
                    const int iR = (int)(q.oi & 0xFFFFu);
                    const int d = hamming_u4(a0, a1, dR + (size_t)iR * 32);
                    if (d < th_high) best = min(best, ((unsigned)d << 16) | (unsigned)iR);
                }
            }
            any_row = __any_sync(0xffffffffu, any_row);
            best = __reduce_min_sync(0xffffffffu, best);
            const int bestDist = (int)(best >> 16), bestIdx = (int)(best & 0xFFFFu);
            if (any_row && range_ok) {
                processed = true;
                if (bestDist < th_high) { out_bd = bestDist; out_bi = bestIdx; }
                if (bestDist < thOrbDist) {
                    if (!classic) {
                        float bestuR = kR[bestIdx].x;
                        float disparity = __fsub_rn(uL, bestuR);
                        if (disparity >= minD && disparity < maxD) {
                            if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                            out_z = __fdiv_rn(prm.bf, disparity);
                            out_u = bestuR;
                            out_dist = bestDist;
                        }
                    } else {
                        // ---- 11x11 SAD slide at the left key's octave (:976-1029) ----
                        const float uR0 = kR[bestIdx].x;
                        const float sf = v.inv_scale[levelL];
                        const float scaleduL = roundf(__fmul_rn(uL, sf));
                        const float scaledvL = roundf(__fmul_rn(vL, sf));
                        const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
                        const float iniu = scaleduR0 + 5.f - 5.f, endu = scaleduR0 + 5.f + 5.f + 1.f;
                        if (!(iniu < 0 || endu >= (float)v.w[levelL])) {
                            const int pitch = v.pitch[levelL];
                            const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
                            const uint8_t* pL = v.pyrL[levelL] + (size_t)f * v.frame_stride[levelL] + (ptrdiff_t)cv * pitch + cu;
                            const uint8_t* pR = v.pyrR[levelL] + (size_t)f * v.frame_stride[levelL] + (ptrdiff_t)cv * pitch + cr;
                            const int cL = __ldg(pL);
                            int acc[11], dlt[11];      // dlt[s] = (right centre at shift s) - (left centre)
#pragma unroll
                            for (int s = 0; s < 11; ++s) { acc[s] = 0; dlt[s] = (int)__ldg(pR + s - 5) - cL; }
                            for (int p = lane; p < 121; p += 32) {
                                const int dy = p / 11 - 5, dx = p - (dy + 5) * 11 - 5;
                                const int lpx = (int)__ldg(pL + dy * pitch + dx);
                                const uint8_t* rrow = pR + dy * pitch + dx - 5;
                                // |(L - cL) - (R - cR_s)| = |L + (cR_s - cL) - R|: IADD3 + IABS + IADD per pixel and shift
#pragma unroll
                                for (int s = 0; s < 11; ++s) acc[s] += abs(lpx + dlt[s] - (int)__ldg(rrow + s));
                            }
                            int bestSad = 0x7fffffff, bestInc = 0;
                            float dists[11];
#pragma unroll
                            for (int s = 0; s < 11; ++s) {
                                const int t = __reduce_add_sync(0xffffffffu, acc[s]);
                                dists[s] = (float)t;
                                if (t < bestSad) { bestSad = t; bestInc = s - 5; }
                            }
                            if (bestInc != -5 && bestInc != 5) {
                                float d1 = 0, d2 = 0, d3 = 0;
#pragma unroll
                                for (int s = 1; s < 10; ++s)
                                    if (s == bestInc + 5) { d1 = dists[s - 1]; d2 = dists[s]; d3 = dists[s + 1]; }
                                const float deltaR = __fdiv_rn(__fsub_rn(d1, d3),
                                    __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                                if (!(deltaR < -1 || deltaR > 1)) {
                                    float bestuR = __fmul_rn(v.scale[levelL],
                                        __fadd_rn(__fadd_rn(scaleduR0, (float)bestInc), deltaR));
                                    float disparity = __fsub_rn(uL, bestuR);
                                    if (disparity >= minD && disparity < maxD) {
                                        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                                        out_z = __fdiv_rn(prm.bf, disparity);
                                        out_u = bestuR;
                                        out_sad = bestSad;
                                        out_dist = bestSad;
                                    }
                                }
                            }
                        }
                    }
                }
            }
        }
        if (lane == 0) {
            uright[iL] = out_u; depth[iL] = out_z;
            o_bd[iL] = out_bd; o_bi[iL] = out_bi; o_sad[iL] = out_sad;
            o_cut[iL] = out_dist;
            if (processed) atomicAdd(&s_processed, 1);
        }
    }
    __syncthreads();
    // ---- frame-level hand-off: the last slice to arrive owns the median cut ----
    if (threadIdx.x == 0) {
        if (s_processed) atomicAdd(&v.proc_count[f], s_processed);
        __threadfence();
        const int ticket = atomicAdd(&v.ticket[f], 1);
        s_last = ticket == kSlices - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int i = threadIdx.x; i < nL; i += kT) {
        const int d = __ldcg(o_cut + i);
        s_dist[i] = d;
        if (d >= 0) atomicAdd(&s_cnt, 1);
    }
    __syncthreads();

    // ---- median cut (:1064-1077 / :1291-1313): drop matches with dist >= 1.5*1.4*median ----
    // only the value of the (n/2)-th smallest distance matters: radix selection (SAD <= 121*255 < 2^16)
    const int n = s_cnt;
    if (n > 0) {
        const int median = select_kth(s_dist, nL, n / 2, s_hist, s_tmp);
        const float c = 1.5f * 1.4f;
        const float thDist = __fmul_rn(c, (float)median);
        for (int i = threadIdx.x; i < nL; i += kT) {
            const int d = s_dist[i];
            if (d >= 0 && !((float)d < thDist)) {
                uright[i] = -1.f; depth[i] = -1.f;
                atomicAdd(&s_removed, 1);
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int processed_all = atomicExch(&v.proc_count[f], 0);     // also re-arms the counters for the next launch
        v.ticket[f] = 0;
        v.nmatch[f] = classic ? (n - s_removed) : (processed_all - s_removed);
    }
}

void launch_stereo(const StereoView& v, const gforb_stereo_params& p, int nframes, int K, cudaStream_t s) {
    const size_t smem = sizeof(RKey) * (size_t)K + sizeof(int) * ((size_t)v.nRows + 2);
    static size_t configured[64] = {};
    if (smem > 48 * 1024) opt_in_smem(stereo_kernel, smem, configured);
    // centre-row window that covers every band: |row - y| < 2*scale + 1  =>  |row - floor(y)| < 2*scale + 2
    float smax = 1.f;
    for (int l = 0; l < kMaxLevels; ++l) smax = v.scale[l] > smax ? v.scale[l] : smax;
    const int band = (int)(2.f * smax) + 3;
    stereo_kernel<<<dim3(nframes, kSlices), kT, smem, s>>>(v, p, K, band);
}

}  // namespace gforb
