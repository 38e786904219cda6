// Stereo association: Frame::ComputeStereoMatches (reference src/Frame.cc:889-1078) and
// Frame::ComputeStereoMatches_Undistorted + PrepareStereoCandidates (src/Frame.cc:1167-1316,
// include/Frame.h:230-263), with ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1768-1784) as __popc.
//
// kSlices CTAs per stereo frame, adjacent in the grid so that they share the frame's keys in L2.  A slice owns the LEFT
// keys of one band of image rows.  The reference builds a per-row table of right keys and walks it per left key; since
// "best" is the lexicographic minimum of (Hamming distance, right index) over the gated candidates (strict '<' while
// walking ascending iR), no table and no sort are needed:
//   phase 0  the slice compacts into shared memory the right keys whose row range [minr, maxr] meets its band
//            (gate record + descriptor, ~1/6 of the frame's keys) and the indices of its own left keys;
//   phase A  ONE THREAD PER LEFT KEY walks the compacted right keys: every lane reads the same shared-memory record
//            (broadcast), tests its own key against it (row, octave +-1, uR window) and keeps a packed minimum -- 32
//            searches in flight per warp and no global load inside the loop;
//   phase B  (classic mode) the 11x11 / 11x21 patches of up to kTile matched keys are staged in shared memory with
//            aligned 32-bit loads, all issued before any is consumed; the 11-shift SAD then runs one thread per
//            (key, shift) out of shared memory, exact in integers (the reference's float sums are integer-valued), and
//            one thread per key fits the parabola in un-fused float ops;
//   phase C  the last slice of a frame to finish (ticket) applies the median cut: only the value of the (n/2)-th
//            smallest distance matters -> radix selection in shared memory.
// INT/LSU-bound.
#include <algorithm>

// Measured and rejected (round 2, late): the SAD of phase B on VABSDIFF4.U8.ACC -- |L + dlt - R| = |(L + da) - (R + db)| with
// da = max(dlt, 0), db = max(-dlt, 0) on funnel-shifted aligned words, four pixels per instruction -- is exact only while
// neither sum leaves a byte.  With that path forced the kernel runs in 0.201 ms (scalar loop: 0.222 ms); with the exact
// guard (largest staged byte of the key + da <= 255, else the scalar loop) the synthetic frames, whose corners saturate,
// fall back so often that both paths run in most warps: 0.248 ms.  An always-exact SIMD form needs a per-byte saturating
// add, which sm_100a emulates in seven instructions (2.75 instructions per pixel against the scalar loop's 4.5).
#include "gforb_internal.cuh"

namespace gforb {

namespace {
#ifndef GFORB_ST_THREADS
#define GFORB_ST_THREADS 256     // re-measured late in round 2 (with the cheaper SAD): 128 threads x 8 CTAs 0.229 ms, 256 x 4 0.219, 256 x 5 0.284, 512 x 2 0.288
#endif
constexpr int kT = GFORB_ST_THREADS;
constexpr int kWarps = kT / 32;
#ifndef GFORB_ST_SLICES
#define GFORB_ST_SLICES 8
#endif
#ifndef GFORB_ST_MINB
#define GFORB_ST_MINB 4
#endif
constexpr int kSlices = GFORB_ST_SLICES;
constexpr int kChunk = 512;          // right keys held in shared memory at a time (a band of a 1000-key frame needs ~170)
constexpr int kTile = 46;            // matched keys whose patches are staged together: 46 * 11 shifts = 506 of 512 SAD tasks
constexpr int kRowWords = 10;        // staged words per patch row: 4 (11 left bytes + alignment) + 6 (21 right bytes + alignment)
constexpr int kKeyWords = 11 * kRowWords;

struct RGate { int rows; uint32_t oi; float x; int pad; };   // rows = minr | maxr << 16, oi = octave << 16 | original index
struct TileKey {                                            // one matched left key of the current SAD tile
    const uint8_t* pl; const uint8_t* pr;                   // word-aligned addresses of the top-left staged words
    float uL, scaleduR0, minD, maxD;
    int iL, level, phL, phR;                                // phX = byte offset of the first patch pixel inside the first staged word
};

__device__ __forceinline__ int hamming_s(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
// row range of a right key in the reference's row table (:906-916 / Frame.h:246-257); the classic reference indexes
// unclamped (keys sit >= 19*scale inside), the restatement clamps, and so does this
__device__ __forceinline__ void row_range(float y, float scale_oct, int nRows, int& minr, int& maxr) {
    const float r = __fmul_rn(2.0f, scale_oct);
    float hi = ceilf(__fadd_rn(y, r)), lo = floorf(__fsub_rn(y, r));
    hi = fminf(hi, (float)(nRows - 1));
    lo = fmaxf(lo, 0.0f);
    minr = (int)lo; maxr = (int)hi;
}
// slice that owns a left key: the band its row (int)vL falls in, out-of-image rows going to the first / last band
__device__ __forceinline__ int slice_of(float vL, int bandRows) {
    const float c = fminf(fmaxf(vL, 0.f), 65535.f);
    return min((int)c / bandRows, kSlices - 1);
}

// k-th smallest (0-based) of the non-negative entries of d[0..n) that are < 2^16, by two 256-bin histogram passes.
// All kT threads call it; hist = 256 ints of shared memory; s_tmp = 2 ints.
__device__ int select_kth(const int* d, int n, int k, int* hist, int* s_tmp) {
    int prefix = 0;
    for (int pass = 0; pass < 2; ++pass) {
        const int shift = pass ? 0 : 8;
        for (int i = threadIdx.x; i < 256; i += kT) hist[i] = 0;
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
stereo_kernel(const StereoView v, const gforb_stereo_params prm, int K) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    // [gate: kChunk x 16 B][rdesc: kChunk x 32 B]   (phase 0 / A)
    //   aliased by [stage: kTile x 110 words][sad: kTile x 11][tkey: kTile]   (phase B)
    //   aliased by [s_dist: K ints]                                          (phase C)
    // [s_best: K][s_left: K ushort][s_match: K ushort]
    RGate* gate = reinterpret_cast<RGate*>(smem_raw);
    uint4* rdesc = reinterpret_cast<uint4*>(smem_raw + sizeof(RGate) * kChunk);
    uint32_t* stage = reinterpret_cast<uint32_t*>(smem_raw);
    int* s_sad = reinterpret_cast<int*>(smem_raw + 4 * kTile * kKeyWords);
    TileKey* tkey = reinterpret_cast<TileKey*>(smem_raw + 4 * kTile * kKeyWords + 4 * kTile * 11 + 8);
    int* s_dist = reinterpret_cast<int*>(smem_raw);
    const size_t shared_area = max((size_t)kChunk * 48, max((size_t)K * 4, (size_t)4 * kTile * kKeyWords + 4 * kTile * 11 + 8 + sizeof(TileKey) * kTile));
    uint32_t* s_best = reinterpret_cast<uint32_t*>(smem_raw + ((shared_area + 15) & ~(size_t)15));
    int* s_bin = reinterpret_cast<int*>(s_best + K);          // bandRows + 2 counters of the left-key counting sort
    uint16_t* s_wlist = reinterpret_cast<uint16_t*>(s_bin + ((v.nRows + kSlices - 1) / kSlices + 2));   // kWarps x kChunk
    uint16_t* s_left = s_wlist + kWarps * kChunk;
    uint16_t* s_match = s_left + K;
    __shared__ int s_hist[256];
    __shared__ int s_tmp[2];
    __shared__ int s_cnt, s_removed, s_processed, s_last, s_nleft, s_nright, s_nmatch;

    const int slice = blockIdx.x, f = v.f0 + blockIdx.y;
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
    const int bandRows = (nRows + kSlices - 1) / kSlices;
    const int bandLo = slice * bandRows, bandHi = min(bandLo + bandRows, nRows) - 1;    // rows a valid left key of this slice can have
    const float minZ = prm.b;
    const float maxD0 = __fdiv_rn(prm.bf, minZ);
    constexpr uint32_t kAny = 0x80000000u;      // s_best bit 31: some right key's row range holds this key's row

    if (threadIdx.x == 0) { s_cnt = 0; s_removed = 0; s_processed = 0; s_nleft = 0; s_nright = 0; s_nmatch = 0; }
    __syncthreads();

    // ---- phase 0a: the left keys of this band, counting-sorted by row so that the 32 keys of a warp span few rows
    for (int i = threadIdx.x; i <= bandRows; i += kT) s_bin[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < nL; i += kT) {
        const float y = kL[i].y;
        if (slice_of(y, bandRows) == slice) atomicAdd(&s_bin[min(max((int)y - bandLo, 0), bandRows - 1) + 1], 1);
    }
    __syncthreads();
    if (threadIdx.x < 32) {                                  // inclusive scan of s_bin[1..bandRows] by one warp
        int carry = 0;
        for (int b0 = 1; b0 <= bandRows; b0 += 32) {
            const int b = b0 + (int)threadIdx.x;
            int x = b <= bandRows ? s_bin[b] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, x, o);
                if ((int)threadIdx.x >= o) x += t;
            }
            if (b <= bandRows) s_bin[b] = x + carry;
            carry += __shfl_sync(0xffffffffu, x, 31);
        }
        if (threadIdx.x == 0) s_nleft = carry;
    }
    __syncthreads();
    // scatter from the back of each bin: afterwards s_bin[b + 1] = first slot of bin b
    for (int i = threadIdx.x; i < nL; i += kT) {
        const float y = kL[i].y;
        if (slice_of(y, bandRows) == slice) {
            const int slot = atomicSub(&s_bin[min(max((int)y - bandLo, 0), bandRows - 1) + 1], 1) - 1;
            s_left[slot] = (uint16_t)i;
            s_best[slot] = ((uint32_t)th_high << 16) | 0xFFFFu;
        }
    }
    __syncthreads();
    const int nleft = s_nleft;

    // ---- phase 0b + A, one chunk of right keys at a time (one chunk in practice)
    for (int r0 = 0; r0 < nR; ) {
        // compact right keys [r0, ...) whose row range meets the band until the chunk is full
        int scanned_to = r0;
        for (; scanned_to < nR; scanned_to += kT) {
            const int held = s_nright;
            __syncthreads();                                       // every thread has read the count before anyone adds to it
            if (held > kChunk - kT) break;
            const int i = scanned_to + threadIdx.x;
            if (i < nR) {
                const gforb_keypoint k = kR[i];
                int minr, maxr;
                row_range(k.y, v.scale[k.octave], nRows, minr, maxr);
                if (maxr >= bandLo && minr <= bandHi) {
                    const int slot = atomicAdd(&s_nright, 1);
                    RGate g;
                    g.rows = minr | (maxr << 16); g.oi = ((uint32_t)k.octave << 16) | (uint32_t)i; g.x = k.x; g.pad = 0;
                    gate[slot] = g;
                    const uint4* src = reinterpret_cast<const uint4*>(dR + (size_t)i * 32);
                    rdesc[2 * slot] = __ldg(src);
                    rdesc[2 * slot + 1] = __ldg(src + 1);
                }
            }
            __syncthreads();
        }
        __syncthreads();
        const int nright = s_nright;
        r0 = scanned_to;

        // one thread per left key of the band; a warp first lists the right keys whose rows meet ITS keys' rows
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        uint16_t* wl = s_wlist + warp * kChunk;
        for (int t0 = warp * 32; t0 < nleft; t0 += kT) {
            const int t = t0 + lane;
            const bool have = t < nleft;
            const int iL = have ? s_left[t] : 0;
            gforb_keypoint kp;
            kp.x = 0.f; kp.y = 0.f; kp.octave = 0;
            if (have) kp = kL[iL];
            const float vL = kp.y, uL = kp.x;
            const int levelL = kp.octave;
            bool ok = classic ? true : !(vL < 0 || vL > (float)(nRows - 1));
            const int row = (int)vL;
            ok = have && ok && row >= 0 && row < nRows;
            const int rlo = __reduce_min_sync(0xffffffffu, ok ? row : 0x7fffffff);
            const int rhi = __reduce_max_sync(0xffffffffu, ok ? row : -1);
            int cnt = 0;
            for (int j0 = 0; j0 < nright; j0 += 32) {
                const int j = j0 + lane;
                bool hit = false;
                if (j < nright) {
                    const int rows = gate[j].rows;
                    hit = (rows >> 16) >= rlo && (rows & 0xFFFF) <= rhi;
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) wl[cnt + __popc(m & ((1u << lane) - 1u))] = (uint16_t)j;
                cnt += __popc(m);
            }
            __syncwarp();
            if (ok) {
                float minD = 0.f, maxD = maxD0;
                if (!classic && v.min_d != nullptr && v.max_d != nullptr) { minD = v.min_d[iL]; maxD = v.max_d[iL]; }
                const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
                const bool range_ok = classic ? !(maxU < 0) : !(maxU < prm.min_x);
                const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32));
                const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32) + 1);
                const uint32_t st = s_best[t];
                uint32_t best = st & ~kAny, any = st & kAny;
                for (int jj = 0; jj < cnt; ++jj) {
                    const int j = wl[jj];                                            // broadcast
                    const RGate q = gate[j];                                         // broadcast
                    const int minr = q.rows & 0xFFFF, maxr = q.rows >> 16;
                    if (row < minr || row > maxr) continue;
                    any = kAny;
                    if (!range_ok) continue;
                    const int oct = (int)(q.oi >> 16);
                    if (oct < levelL - 1 || oct > levelL + 1) continue;
                    if (q.x >= minU && q.x <= maxU) {
                        const int d = hamming_s(a0, a1, rdesc[2 * j], rdesc[2 * j + 1]);
                        if (d < th_high) best = min(best, ((uint32_t)d << 16) | (q.oi & 0xFFFFu));
                    }
                }
                s_best[t] = best | any;
            }
            __syncwarp();
        }
        __syncthreads();
        if (threadIdx.x == 0) s_nright = 0;
        __syncthreads();
    }

    // ---- results of the search; classic mode queues the keys that go on to the SAD refinement
    for (int t = threadIdx.x; t < nleft; t += kT) {
        const int iL = s_left[t];
        const gforb_keypoint kp = kL[iL];
        const float vL = kp.y, uL = kp.x;
        bool ok = classic ? true : !(vL < 0 || vL > (float)(nRows - 1));
        const int row = (int)vL;
        ok = ok && row >= 0 && row < nRows;
        float minD = 0.f, maxD = maxD0;
        if (!classic && v.min_d != nullptr && v.max_d != nullptr) { minD = v.min_d[iL]; maxD = v.max_d[iL]; }
        const float maxU = __fsub_rn(uL, minD);
        const bool range_ok = classic ? !(maxU < 0) : !(maxU < prm.min_x);
        const uint32_t st = s_best[t];
        const int bestDist = (int)((st & ~kAny) >> 16), bestIdx = (int)(st & 0xFFFFu);
        float out_u = -1.f, out_z = -1.f;
        int out_bd = th_high, out_bi = -1, out_dist = -1;
        if (ok && (st & kAny) && range_ok) {
            atomicAdd(&s_processed, 1);
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
                    s_match[atomicAdd(&s_nmatch, 1)] = (uint16_t)t;
                }
            }
        }
        uright[iL] = out_u; depth[iL] = out_z;
        o_bd[iL] = out_bd; o_bi[iL] = out_bi; o_sad[iL] = -1;
        o_cut[iL] = out_dist;
    }
    __syncthreads();

    // ---- phase B: 11x11 SAD slide at the left key's octave (:976-1029), kTile keys at a time
    const int nmatch = classic ? s_nmatch : 0;
    for (int m0 = 0; m0 < nmatch; m0 += kTile) {
        const int nt = min(kTile, nmatch - m0);
        if (threadIdx.x < nt) {
            const int t = s_match[m0 + threadIdx.x];
            const int iL = s_left[t];
            const gforb_keypoint kp = kL[iL];
            const int bestIdx = (int)(s_best[t] & 0xFFFFu);
            const int levelL = kp.octave;
            const float uR0 = kR[bestIdx].x;
            const float sf = v.inv_scale[levelL];
            const float scaleduL = roundf(__fmul_rn(kp.x, sf));
            const float scaledvL = roundf(__fmul_rn(kp.y, sf));
            const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
            const float iniu = scaleduR0 + 5.f - 5.f, endu = scaleduR0 + 5.f + 5.f + 1.f;
            TileKey tk;
            tk.iL = iL; tk.level = levelL; tk.uL = kp.x; tk.scaleduR0 = scaleduR0; tk.minD = 0.f; tk.maxD = maxD0;
            tk.pl = nullptr; tk.pr = nullptr; tk.phL = 0; tk.phR = 0;
            if (!(iniu < 0 || endu >= (float)v.w[levelL])) {
                const int pitch = v.pitch[levelL];
                const int cu = (int)scaleduL, cv = (int)scaledvL, cr = (int)scaleduR0;
                const uint8_t* pL = v.pyrL[levelL] + (size_t)f * v.frame_stride[levelL] + (ptrdiff_t)(cv - 5) * pitch + (cu - 5);
                const uint8_t* pR = v.pyrR[levelL] + (size_t)f * v.frame_stride[levelL] + (ptrdiff_t)(cv - 5) * pitch + (cr - 10);
                tk.phL = (int)((uintptr_t)pL & 3); tk.phR = (int)((uintptr_t)pR & 3);
                tk.pl = pL - tk.phL; tk.pr = pR - tk.phR;
            }
            tkey[threadIdx.x] = tk;
        }
        __syncthreads();
        // stage: one thread per (key, patch row) issues that row's 4 + 6 aligned words; nothing waits on anything
        for (int idx = threadIdx.x; idx < nt * 11; idx += kT) {
            const int k = idx / 11, rowi = idx - k * 11;
            const TileKey& tk = tkey[k];
            if (tk.pl == nullptr) continue;
            const ptrdiff_t ro = (ptrdiff_t)rowi * v.pitch[tk.level];
            const uint32_t* gl = reinterpret_cast<const uint32_t*>(tk.pl + ro);
            const uint32_t* gr = reinterpret_cast<const uint32_t*>(tk.pr + ro);
            uint32_t w[kRowWords];
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = __ldg(gl + i);
#pragma unroll
            for (int i = 0; i < 6; ++i) w[4 + i] = __ldg(gr + i);
            uint32_t* dst = stage + k * kKeyWords + rowi * kRowWords;
#pragma unroll
            for (int i = 0; i < kRowWords; ++i) dst[i] = w[i];
        }
        __syncthreads();
        // one thread per (key, shift): |(L - cL) - (R_s - cR_s)| = |L + (cR_s - cL) - R_s| summed over the 11x11 window
        for (int idx = threadIdx.x; idx < nt * 11; idx += kT) {
            const int k = idx / 11, sft = idx - k * 11;
            const TileKey& tk = tkey[k];
            int acc = 0;
            if (tk.pl != nullptr) {
                const uint8_t* base = reinterpret_cast<const uint8_t*>(stage + k * kKeyWords);
                const uint8_t* lb = base + tk.phL;
                const uint8_t* rb = base + 16 + tk.phR + sft;
                const int dlt = (int)rb[5 * 4 * kRowWords + 5] - (int)lb[5 * 4 * kRowWords + 5];
#pragma unroll
                for (int dy = 0; dy < 11; ++dy) {
#pragma unroll
                    for (int dx = 0; dx < 11; ++dx)
                        acc = (int)__sad((int)lb[dy * 4 * kRowWords + dx] + dlt, (int)rb[dy * 4 * kRowWords + dx], (unsigned)acc);   // |a - b| + c: one VABSDIFF
                }
            }
            s_sad[idx] = acc;
        }
        __syncthreads();
        if (threadIdx.x < nt) {
            const TileKey& tk = tkey[threadIdx.x];
            if (tk.pl != nullptr) {
                const int* sd = s_sad + threadIdx.x * 11;
                int bestSad = 0x7fffffff, bestInc = 0;
#pragma unroll
                for (int s2 = 0; s2 < 11; ++s2) {
                    const int t2 = sd[s2];
                    if (t2 < bestSad) { bestSad = t2; bestInc = s2 - 5; }
                }
                if (bestInc != -5 && bestInc != 5) {
                    const float d1 = (float)sd[bestInc + 4], d2 = (float)sd[bestInc + 5], d3 = (float)sd[bestInc + 6];
                    const float deltaR = __fdiv_rn(__fsub_rn(d1, d3),
                        __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                    if (!(deltaR < -1 || deltaR > 1)) {
                        float bestuR = __fmul_rn(v.scale[tk.level], __fadd_rn(__fadd_rn(tk.scaleduR0, (float)bestInc), deltaR));
                        float disparity = __fsub_rn(tk.uL, bestuR);
                        if (disparity >= tk.minD && disparity < tk.maxD) {
                            if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)tk.uL - 0.01); }
                            depth[tk.iL] = __fdiv_rn(prm.bf, disparity);
                            uright[tk.iL] = bestuR;
                            o_sad[tk.iL] = bestSad;
                            o_cut[tk.iL] = bestSad;
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

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
    // only the value of the (n/2)-th smallest distance matters: radix selection (SAD <= 121*510 < 2^16)
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

size_t stereo_smem_bytes(int K, int nRows) {
    size_t shared_area = (size_t)kChunk * 48;
    shared_area = std::max(shared_area, (size_t)K * 4);
    shared_area = std::max(shared_area, (size_t)4 * kTile * kKeyWords + 4 * kTile * 11 + 8 + sizeof(TileKey) * kTile);
    return ((shared_area + 15) & ~(size_t)15) + (size_t)K * 4 + 4 * ((size_t)(nRows + kSlices - 1) / kSlices + 2) +
           (size_t)kWarps * kChunk * 2 + (size_t)K * 2 * 2;
}

void launch_stereo(const StereoView& v, const gforb_stereo_params& p, int nframes, int K, cudaStream_t s) {
    const size_t smem = stereo_smem_bytes(K, v.nRows);
    static size_t configured[64] = {};
    // the 48 KB a launch gets without the opt-in covers static + dynamic shared memory together, and the kernel holds ~1 KB of
    // static: a dynamic size just under 48 KB (K around 2,000 keys) failed to launch -- opt in from 32 KB on
    if (smem > 32 * 1024) opt_in_smem(stereo_kernel, smem, configured);
    stereo_kernel<<<dim3(kSlices, nframes), kT, smem, s>>>(v, p, K);
}

}  // namespace gforb
