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
constexpr int kSlices = 8;
struct RKey { short minr, maxr; int octave; float x; };

__device__ __forceinline__ int hamming_u4(const uint4 a0, const uint4 a1, const uint8_t* b) {
    const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(b));
    const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(b) + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
}  // namespace

__global__ void __launch_bounds__(kT)
stereo_kernel(const StereoView v, const gforb_stereo_params prm, int K) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    RKey* rk = reinterpret_cast<RKey*>(smem_raw);
    int* s_dist = reinterpret_cast<int*>(smem_raw + sizeof(RKey) * (size_t)K);
    __shared__ int s_cnt, s_median, s_removed, s_processed, s_last;

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

    // right keys: row band of radius 2*scale[octave] (:906-916 / Frame.h:246-257)
    for (int i = threadIdx.x; i < nR; i += kT) {
        const gforb_keypoint k = kR[i];
        const float r = __fmul_rn(2.0f, v.scale[k.octave]);
        float hi = ceilf(__fadd_rn(k.y, r)), lo = floorf(__fsub_rn(k.y, r));
        hi = fminf(hi, (float)(nRows - 1));   // classic reference indexes unclamped (keys sit >= 19*scale inside)
        lo = fmaxf(lo, 0.0f);
        RKey q;
        q.minr = (short)(int)lo; q.maxr = (short)(int)hi; q.octave = k.octave; q.x = k.x;
        rk[i] = q;
    }
    if (threadIdx.x == 0) { s_cnt = 0; s_removed = 0; s_processed = 0; }
    __syncthreads();

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
            for (int iR = lane; iR < nR; iR += 32) {
                const RKey q = rk[iR];
                if (row < q.minr || row > q.maxr) continue;
                any_row = true;
                if (!range_ok) continue;
                if (q.octave < levelL - 1 || q.octave > levelL + 1) continue;
                if (q.x >= minU && q.x <= maxU) {
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
                            int acc[11];
#pragma unroll
                            for (int s = 0; s < 11; ++s) acc[s] = 0;
                            for (int p = lane; p < 121; p += 32) {
                                const int dy = p / 11 - 5, dx = p - (dy + 5) * 11 - 5;
                                const int lv = (int)__ldg(pL + dy * pitch + dx) - cL;
                                const uint8_t* rrow = pR + dy * pitch + dx;
#pragma unroll
                                for (int s = 0; s < 11; ++s) {
                                    const int inc = s - 5;
                                    const int rv = (int)__ldg(rrow + inc) - (int)__ldg(pR + inc);
                                    acc[s] += abs(lv - rv);
                                }
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
    const int n = s_cnt;
    if (n > 0) {
        const int m = n / 2;
        for (int i = threadIdx.x; i < nL; i += kT) {
            const int d = s_dist[i];
            if (d < 0) continue;
            int lt = 0, le = 0;
            for (int j = 0; j < nL; ++j) {
                const int e = s_dist[j];
                if (e < 0) continue;
                lt += e < d;
                le += e <= d;
            }
            if (lt <= m && m < le) s_median = d;   // every writer holds the same value
        }
        __syncthreads();
        const float c = 1.5f * 1.4f;
        const float thDist = __fmul_rn(c, (float)s_median);
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
    const size_t smem = (sizeof(RKey) + sizeof(int)) * (size_t)K;
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        cudaFuncSetAttribute(stereo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = smem;
    }
    stereo_kernel<<<dim3(nframes, kSlices), kT, smem, s>>>(v, p, K);
}

}  // namespace gforb
