// Brute-force 256-bit Hamming matcher with best / second-best selection and ratio test
// (BASELINE.json configs[3]; SURVEY.md 8b "Hamming" row).
//
// Reference semantics: the nearest-neighbour loop every ORBmatcher search shares, here over ALL train
// descriptors (reference src/ORBmatcher.cc:555-577 in SearchForInitialization, :174-206 SearchByProjection):
//     if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx = j; } else if (dist < bestDist2) bestDist2 = dist;
// followed by `bestDist <= TH && bestDist < (float)bestDist2 * mfNNratio` (:579-581).  The loop result does
// not depend on visiting order except for the index of equal minima (first j wins), so it is restated as
//     best   = min over j of key_j,  key_j = dist_j << 22 | j          (lowest distance, then lowest index)
//     second = min(second, max(best_before, key_j))                     (second smallest of the multiset)
// Distance = __popc over the 8 words of a ^ b (ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1768-1784).
//
// Mapping: one query descriptor per thread (8 registers); train descriptors stream through shared memory in
// tiles read as broadcast LDS.128; the train set is cut into `nseg` segments (grid.y) so that small query sets
// still fill 148 SMs; a second kernel merges the per-segment (best, second) pairs and applies the tests.
// Bound: POPC issue (16 lanes/clk/SM measured, tools/int_peaks.cu): 8 POPC per pair -> 2 pairs/clk/SM.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kQ = 128;          // queries per CTA (one per thread)
constexpr int kTile = 256;       // train descriptors per shared tile (8 KB)
constexpr int kIdxBits = 22;     // train index bits inside a key: nB <= 4,194,304
constexpr uint32_t kNoKey = 0xFFFFFFFFu;

__device__ __forceinline__ int hamming8(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return (__popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z)) +
           (__popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y)) +
           (__popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w));
}
}  // namespace

// part: [nA][nseg] uint2 {best key, second key}
__global__ void __launch_bounds__(kQ)
match_scan_kernel(const uint4* __restrict__ A, int nA, const uint4* __restrict__ B, int nB, int seg_len,
                  uint2* __restrict__ part) {
    __shared__ uint4 tile[kTile * 2];
    const int q = blockIdx.x * kQ + threadIdx.x;
    const int seg = blockIdx.y, nseg = gridDim.y;
    const int j0 = seg * seg_len, j1 = min(j0 + seg_len, nB);
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (q < nA) { a0 = __ldg(A + 2 * (size_t)q); a1 = __ldg(A + 2 * (size_t)q + 1); }
    uint32_t best = kNoKey, second = kNoKey;
    for (int t0 = j0; t0 < j1; t0 += kTile) {
        const int n = min(kTile, j1 - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * n; i += kQ) tile[i] = __ldg(B + 2 * (size_t)t0 + i);
        __syncthreads();
        uint32_t kbase = (uint32_t)t0;
#pragma unroll 4
        for (int j = 0; j < n; ++j) {
            const int d = hamming8(a0, a1, tile[2 * j], tile[2 * j + 1]);
            const uint32_t key = ((uint32_t)d << kIdxBits) + kbase + (uint32_t)j;
            second = min(second, max(best, key));
            best = min(best, key);
        }
    }
    if (q < nA) part[(size_t)q * nseg + seg] = make_uint2(best, second);
}

__global__ void match_merge_kernel(const uint2* __restrict__ part, int nA, int nseg, float ratio, int threshold,
                                   int32_t* __restrict__ best_idx, int32_t* __restrict__ best_dist,
                                   int32_t* __restrict__ second_dist, int32_t* __restrict__ match) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nA) return;
    uint32_t best = kNoKey, second = kNoKey;
    for (int s = 0; s < nseg; ++s) {
        const uint2 p = part[(size_t)q * nseg + s];
        second = min(min(second, p.y), max(best, p.x));
        best = min(best, p.x);
    }
    const int bd = best == kNoKey ? INT32_MAX : (int)(best >> kIdxBits);
    const int sd = second == kNoKey ? INT32_MAX : (int)(second >> kIdxBits);
    const int bi = best == kNoKey ? -1 : (int)(best & ((1u << kIdxBits) - 1));
    if (best_idx) best_idx[q] = bi;
    if (best_dist) best_dist[q] = bd;
    if (second_dist) second_dist[q] = sd;
    // src/ORBmatcher.cc:579-581: int -> float conversion of bestDist2, float multiply, int promoted to float for '<'
    if (match) match[q] = (bi >= 0 && bd <= threshold && (float)bd < __fmul_rn((float)sd, ratio)) ? bi : -1;
}

int match_segments(int nA, int nB, int sm_count) {
    const int qblocks = (nA + kQ - 1) / kQ;
    int nseg = (4 * sm_count + qblocks - 1) / std::max(qblocks, 1);          // aim at >= 4 CTAs per SM
    nseg = std::min(nseg, (nB + 2 * kTile - 1) / (2 * kTile));                 // >= 2 tiles per segment
    return std::max(nseg, 1);
}

size_t match_scratch_bytes(int nA, int nseg) { return (size_t)std::max(nA, 1) * nseg * sizeof(uint2); }

int match_max_train() { return 1 << kIdxBits; }

void launch_match(const uint8_t* dA, int nA, const uint8_t* dB, int nB, float ratio, int threshold, int nseg,
                  void* scratch, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match,
                  cudaStream_t s) {
    if (nA <= 0) return;
    int seg_len = (nB + nseg - 1) / nseg;
    seg_len = std::max((seg_len + kTile - 1) / kTile * kTile, kTile);
    dim3 grd((nA + kQ - 1) / kQ, nseg);
    match_scan_kernel<<<grd, kQ, 0, s>>>(reinterpret_cast<const uint4*>(dA), nA, reinterpret_cast<const uint4*>(dB), nB,
                                         seg_len, reinterpret_cast<uint2*>(scratch));
    match_merge_kernel<<<(nA + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(scratch), nA, nseg, ratio, threshold,
                                                        best_idx, best_dist, second_dist, match);
}

// the merge + acceptance tests alone: the tensor-core path (k_match_tc.cu) writes the same per-segment key pairs
void launch_match_merge(const void* part, int nA, int nseg, float ratio, int threshold, int32_t* best_idx, int32_t* best_dist,
                        int32_t* second_dist, int32_t* match, cudaStream_t s) {
    if (nA <= 0) return;
    match_merge_kernel<<<(nA + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(part), nA, nseg, ratio, threshold,
                                                        best_idx, best_dist, second_dist, match);
}

// ------------------------------------------------------------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:397-458), the descriptor part: per map point, the
// N observed descriptors' pairwise distances, per row the median sorted[int(0.5 * (N - 1))] (the row includes its own
// 0), and the first row with the least median.  One warp per map point, the set staged in shared memory; lane = row.
// A row's median is found without sorting: distances are 0..256, so the k-th smallest is the least v with
// #{j : d_ij <= v} > k -- 9 bisection steps, each re-scanning the row against broadcast shared-memory reads.
namespace {
constexpr int kDdWarps = 4;
}
__global__ void __launch_bounds__(kDdWarps * 32)
distinctive_kernel(const uint4* __restrict__ desc, const int32_t* __restrict__ set_start, int npoints, int max_n,
                   int32_t* __restrict__ best_idx, int32_t* __restrict__ best_median) {
    extern __shared__ uint4 dd_tile[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int p = blockIdx.x * kDdWarps + wid;
    if (p >= npoints) return;
    const int s0 = set_start[p], n = set_start[p + 1] - s0;
    if (n <= 0) { if (lane == 0) { best_idx[p] = -1; best_median[p] = INT32_MAX; } return; }
    uint4* t = dd_tile + (size_t)wid * max_n * 2;
    for (int i = lane; i < 2 * n; i += 32) t[i] = __ldg(desc + 2 * (size_t)s0 + i);
    __syncwarp();
    const int k = (int)(0.5 * (n - 1));
    uint32_t best = kNoKey;                        // median << 16 | row: least median, then first row
    for (int i = lane; i < n; i += 32) {
        const uint4 a0 = t[2 * i], a1 = t[2 * i + 1];
        int lo = 0, hi = 256;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int c = 0;
            for (int j = 0; j < n; ++j) c += hamming8(a0, a1, t[2 * j], t[2 * j + 1]) <= mid ? 1 : 0;
            if (c > k) hi = mid; else lo = mid + 1;
        }
        best = min(best, ((uint32_t)lo << 16) | (uint32_t)i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
    if (lane == 0) { best_idx[p] = (int)(best & 0xFFFFu); best_median[p] = (int)(best >> 16); }
}

int distinctive_max_set() { return 1024; }          // rows per set: 4 warps x 1024 x 32 B = 128 KB of shared memory

int launch_distinctive(const uint8_t* desc, const int32_t* set_start, int npoints, int max_n, int32_t* best_idx,
                       int32_t* best_median, cudaStream_t s) {
    if (npoints <= 0) return 0;
    const size_t smem = (size_t)kDdWarps * std::max(max_n, 1) * 32;
    static size_t have[64] = {};
    if (smem > 32 * 1024 && opt_in_smem(distinctive_kernel, smem, have)) return -1;      // static + dynamic share the 48 KB default
    distinctive_kernel<<<(npoints + kDdWarps - 1) / kDdWarps, kDdWarps * 32, smem, s>>>(
        reinterpret_cast<const uint4*>(desc), set_start, npoints, std::max(max_n, 1), best_idx, best_median);
    return 0;
}

}  // namespace gforb
