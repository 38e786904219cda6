// Brute-force 256-bit Hamming matcher as a tensor-core contraction (BASELINE.json configs[3]).
//
// popc(a ^ b) over 256 bits is a dot product in disguise: with every bit expanded to an int8 in {-1, +1},
//     a . b = (#equal bits) - (#different bits) = 256 - 2 * hamming(a, b),
// an EXACT K = 256 integer contraction -- the one op of this library that the 5th-generation tensor cores can take
// (tcgen05.mma kind::i8, int32 accumulators in TMEM).  The POPC kernel of k_match.cu sits at 91 % of the POPC issue rate
// (8 POPC per pair); here a pair costs one int8 MAC lane and the SM's integer pipes only run the best / second-best scan.
//
// One CTA = 128 query rows x a segment of the train set, walked in tiles of 256 train rows:
//   warp 0      TMA producer: the CTA's A block once (2 x [128 rows x 128 B], SWIZZLE_128B), then per tile the B block
//               (2 x [256 rows x 128 B]) into a 2-stage ring, completing on mbarriers;
//   warp 1      MMA issuer (one elected lane): 8 x tcgen05.mma (M 128, N 256, K 32) per tile into one of two 256-column
//               TMEM accumulator stages; tcgen05.commit releases the smem stage and publishes the accumulator;
//   warps 2-9   epilogue: thread = query row = TMEM lane, two warps per lane group each scanning one half of the tile's
//               columns; tcgen05.ld 32 columns at a time; eight columns are screened against the running second-best with
//               four 3-input max, and only groups that hold a candidate run the (best, second) min / max recurrence of
//               k_match.cu on keys distance << 22 | train index.
// The per-segment (best, second) pairs go through the same match_merge_kernel as the POPC path: identical outputs.
// Measured on B200, 64k x 64k (round 2): 0.81 ms = 5.3 T pairs/s = 2.7 P int8 op/s (POPC kernel: 8.07 ms), the tensor pipe
// 48 % busy in the ncu capture (profiles/ncu_r2_match_tc.json).  How it got there: scanning every column, the epilogue's ALU pipe was the
// bound (four epilogue warps 1.37 ms, eight 1.10 ms, sixteen 1.16 ms); the threshold screen above took it to 0.85 ms and the
// wave-aware segment count below to 0.81 ms.  Tried and not kept: two columns per register as packed 16-bit keys scanned
// with VIMNMX.U16x2 (1.92 ms: the packed form issues at a fraction of the 32-bit rate); tcgen05.ld of the next 32 columns in
// flight during the screen (0.86 ms, 168 registers); 256 query rows per CTA against 128-row train tiles in a 4-stage ring,
// i.e. half the L2 -> shared-memory traffic (0.82 ms: that traffic, 36 % of the L2 peak, is not what the kernel waits for).
#include <climits>
#include "gforb_internal.cuh"
#include "tma.cuh"

namespace gforb {

namespace {
constexpr int kBM = 128;             // query rows per CTA (= TMEM lanes)
constexpr int kBN = 256;             // train rows per tile (= TMEM columns of one accumulator stage)
constexpr int kRowBytes = 256;       // one expanded descriptor: 256 int8
constexpr int kHalf = 128;           // bytes of K per TMA box / swizzle atom
constexpr int kStages = 2;           // B ring
constexpr int kIdxBitsTc = 22;       // as k_match.cu: train index bits inside a global key
constexpr uint32_t kNoKeyTc = 0xFFFFFFFFu;
constexpr uint32_t kABytes = kBM * kRowBytes;       // 32 KB
constexpr uint32_t kBBytes = kBN * kRowBytes;       // 64 KB per stage
#ifndef GFORB_TC_EPI_WARPS
#define GFORB_TC_EPI_WARPS 8
#endif
constexpr int kEpiWarps = GFORB_TC_EPI_WARPS;      // kColSplit per TMEM lane group: each takes one slice of a tile's columns
constexpr int kColSplit = kEpiWarps / 4;
constexpr int kColsPerWarp = kBN / kColSplit;
constexpr int kTcThreads = 64 + 32 * kEpiWarps;

// ---- PTX wrappers (sm_100a) ----
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {      // arrives on `bar` when all earlier tcgen05.mma of this thread are done
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, int8 x int8 -> int32, M 128 x N 256 x K 32
__device__ __forceinline__ void tc_mma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// shared-memory matrix descriptor, K-major, SWIZZLE_128B: rows of 128 bytes, 8-row groups 1,024 bytes apart
// (cute/arch/mma_sm100_desc.hpp SmemDescriptor: start >> 4 [0,14), LBO [16,30) unused here, SBO >> 4 [32,46), version 1 [46,48),
// layout type 2 = SWIZZLE_128B [61,64))
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (InstrDescriptor): D = S32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);

#define GFORB_TMEM_LD32(taddr, r)                                                                                           \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                  \
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                  \
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                  \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),          \
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),    \
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),  \
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])   \
                 : "r"(taddr))
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
}  // namespace

// descriptor bits -> int8 {-1, +1}: out row i = 256 bytes, byte k = bit k of descriptor i (byte k / 8, bit k % 8)
__global__ void __launch_bounds__(256)
match_expand_kernel(const uint8_t* __restrict__ desc, int n, int8_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;          // one thread = one descriptor byte = 8 output bytes
    if (i >= n * 32) return;
    const uint32_t x = desc[i];
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        lo |= ((x >> b) & 1u ? 0x01u : 0xFFu) << (8 * b);
        hi |= ((x >> (b + 4)) & 1u ? 0x01u : 0xFFu) << (8 * b);
    }
    reinterpret_cast<uint2*>(out)[i] = make_uint2(lo, hi);
}

// part: [nA][nseg][kColSplit column slices] uint2 {best key, second key}, keys = distance << 22 | train index (as match_scan_kernel writes them)
__global__ void __launch_bounds__(kTcThreads, 1)
match_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int nA, int nB,
                int tiles_per_seg, uint2* __restrict__ part) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t* base = reinterpret_cast<uint8_t*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);      // SWIZZLE_128B atoms: 1,024-byte aligned
    uint8_t* sA = base;                                      // [2][128 x 128 B]
    uint8_t* sB = base + kABytes;                            // [kStages][2][256 x 128 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kABytes + kStages * kBBytes);
    uint64_t* bar_a = bars;
    uint64_t* full = bars + 1;                               // [kStages] TMA -> MMA
    uint64_t* empty = bars + 1 + kStages;                    // [kStages] MMA -> TMA
    uint64_t* tfull = bars + 1 + 2 * kStages;                // [2] MMA -> epilogue
    uint64_t* tempty = bars + 3 + 2 * kStages;               // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5 + 2 * kStages);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seg = blockIdx.y, nseg = gridDim.y;
    const int ntiles = (nB + kBN - 1) / kBN;
    const int t0 = seg * tiles_per_seg, t1 = min(t0 + tiles_per_seg, ntiles);
    const int row0 = blockIdx.x * kBM;

    if (threadIdx.x == 0) {
        mbar_init(bar_a, 1);
        for (int s = 0; s < kStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(tfull + s, 1); mbar_init(tempty + s, kEpiWarps); }
    }
    if (warp == 1) {                                         // TMEM: both accumulator stages = all 512 columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0 && t0 < t1) {
            mbar_expect_tx(bar_a, kABytes);
            tma_load_2d(sA, &mapA, 0, row0, bar_a);
            tma_load_2d(sA + kABytes / 2, &mapA, kHalf, row0, bar_a);
            for (int t = t0; t < t1; ++t) {
                const int j = t - t0, s = j % kStages;
                mbar_wait(empty + s, ((j / kStages) & 1) ^ 1);
                mbar_expect_tx(full + s, kBBytes);
                uint8_t* dst = sB + (size_t)s * kBBytes;
                tma_load_2d(dst, &mapB, 0, t * kBN, full + s);
                tma_load_2d(dst + kBBytes / 2, &mapB, kHalf, t * kBN, full + s);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0 && t0 < t1) {
            mbar_wait(bar_a, 0);
            for (int t = t0; t < t1; ++t) {
                const int j = t - t0, s = j % kStages, as = j & 1;
                mbar_wait(tempty + as, ((j >> 1) & 1) ^ 1);          // the epilogue has drained this accumulator stage
                mbar_wait(full + s, (j / kStages) & 1);               // the tile's B block has landed
                tc_fence_after();
                const uint32_t d = tmem + (uint32_t)(as * kBN);
                const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB + (size_t)s * kBBytes);
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int k = 0; k < 4; ++k) {                     // K = 32 bytes per instruction inside the 128-byte atom
                        const uint64_t da = tc_smem_desc(a_addr + h * (kABytes / 2) + k * 32);
                        const uint64_t db = tc_smem_desc(b_addr + h * (kBBytes / 2) + k * 32);
                        tc_mma_i8(d, da, db, kIdesc, (h | k) ? 1u : 0u);
                    }
                tc_commit(empty + s);                                 // smem stage free once these MMAs have read it
                tc_commit(tfull + as);                                // accumulator complete
            }
        }
    } else {
        // ===== epilogue: thread = query row; warp w may touch TMEM lanes 32 * (w % 4) .. + 31 =====
        const int lg = warp & 3, ch = (warp - 2) >> 2;                // lane group; column half of every tile this warp scans
        const int row = row0 + lg * 32 + lane;
        uint32_t best = kNoKeyTc, second = kNoKeyTc;                  // global keys: distance << 22 | train index
        // A column can only change (best, second) if its key is below `second`.  This thread meets its columns in rising
        // index order, so a column whose distance merely EQUALS the second's cannot (its index is larger): the test is
        // dot > thr with thr = 256 - 2 * distance(second).  Eight columns are screened with four 3-input max; the full
        // key arithmetic runs only for the groups that hold a candidate -- after the first tiles, a few per cent of them.
        int thr = -100000;                                            // nothing seen yet: every group is scanned
        for (int t = t0; t < t1; ++t) {
            const int j = t - t0, as = j & 1;
            mbar_wait(tfull + as, (j >> 1) & 1);
            tc_fence_after();
            const int valid = min(kBN, nB - t * kBN);                 // train rows of this tile that exist
            const uint32_t gbase = (uint32_t)(t * kBN);
            const uint32_t taddr = tmem + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * kBN);
            auto scan32 = [&](const uint32_t (&r)[32], int c0) {
#pragma unroll
                for (int g = 0; g < 32; g += 8) {
                    int m = __vimax3_s32((int)r[g], (int)r[g + 1], (int)r[g + 2]);
                    m = __vimax3_s32(m, (int)r[g + 3], (int)r[g + 4]);
                    m = __vimax3_s32(m, (int)r[g + 5], (int)r[g + 6]);
                    m = max(m, (int)r[g + 7]);
                    if (m > thr) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int col = c0 + g + i;
                            if (col < valid) {
                                // distance << 22 | index = (256 - dot) << 21 | index   (256 - dot is even)
                                const uint32_t key = ((uint32_t)(256 - (int)r[g + i]) << (kIdxBitsTc - 1)) + gbase + (uint32_t)col;
                                second = min(second, max(best, key));
                                best = min(best, key);
                            }
                        }
                        thr = second == kNoKeyTc ? -100000 : 256 - 2 * (int)(second >> kIdxBitsTc);
                    }
                }
            };
#pragma unroll 1
            for (int c0 = ch * kColsPerWarp; c0 < (ch + 1) * kColsPerWarp; c0 += 32) {
                uint32_t r[32];
                GFORB_TMEM_LD32(taddr + (uint32_t)c0, r);
                tc_wait_ld();
                scan32(r, c0);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + as);                  // this warp's lanes of the stage are read
        }
        if (row < nA) part[((size_t)row * nseg + seg) * kColSplit + ch] = make_uint2(best, second);     // the two halves merge like two segments
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

namespace {
// 2-D u8 tensor map over the expanded descriptors: {256 bytes, rows}, box {128 bytes, box_rows}, SWIZZLE_128B
int encode_desc_map(CUtensorMap* map, const void* base, uint64_t rows, uint32_t box_rows) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return -1;
        fn = (EncodeFn)p;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)kRowBytes, rows};
    const cuuint64_t strides[1] = {(cuuint64_t)kRowBytes};
    const cuuint32_t box[2] = {(cuuint32_t)kHalf, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return (int)fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}
size_t tc_smem_bytes() { return 1024 + kABytes + (size_t)kStages * kBBytes + 16 * 8; }
}  // namespace

// the tensor-core path pays from a few thousand descriptors a side (expansion + 1 CTA per SM); below that the POPC kernel wins
bool match_tc_eligible(int nA, int nB) {
    static const bool off = getenv("GFORB_NO_TC") != nullptr;          // A/B switch for measurements and parity tests
    return !off && nA >= 2048 && nB >= 2048;
}
int match_tc_segments(int nA, int nB, int sm_count) {
    // One CTA per SM at a time (161 KB of shared memory), so the grid runs in waves of sm_count CTAs, and every CTA pays a
    // start-up of about a dozen tile times: pipeline fill, and -- the larger part -- its first tiles run the epilogue's
    // slow path until the running second-best has settled (measured: 64k x 64k 0.834 ms with 1 segment, 0.809 with 4;
    // 32k x 32k 0.279 ms with 2, 0.295 with 4).  Cost of n segments = waves x (tiles per CTA + 12); a tie goes to fewer segments.
    const int mblocks = (nA + kBM - 1) / kBM, ntiles = (nB + kBN - 1) / kBN;
    const int cap = std::max(std::min(ntiles / 4, 64), 1);
    int best = 1;
    long best_cost = LONG_MAX;
    for (int n = 1; n <= cap; ++n) {
        const int per = (ntiles + n - 1) / n, used = (ntiles + per - 1) / per;         // segments that get tiles
        const long ctas = (long)mblocks * used, waves = (ctas + sm_count - 1) / sm_count;
        const long cost = waves * (per + 12);
        if (cost < best_cost) { best_cost = cost; best = n; }
    }
    return best;
}
size_t match_tc_scratch_bytes(int nA, int nB, int nseg) {
    const size_t pa = (size_t)((nA + kBM - 1) / kBM) * kBM, pb = (size_t)((nB + kBN - 1) / kBN) * kBN;
    return (((size_t)nA * nseg * kColSplit * sizeof(uint2) + 1023) & ~(size_t)1023) + (pa + pb) * kRowBytes;
}

// returns 0, or -1 when the tensor maps / shared-memory opt-in cannot be set up (the caller falls back to the POPC kernels)
int launch_match_tc(const uint8_t* dA, int nA, const uint8_t* dB, int nB, float ratio, int threshold, int nseg, void* scratch,
                    int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match, cudaStream_t s) {
    const size_t pa = (size_t)((nA + kBM - 1) / kBM) * kBM, pb = (size_t)((nB + kBN - 1) / kBN) * kBN;
    uint8_t* base = static_cast<uint8_t*>(scratch);
    uint2* part = reinterpret_cast<uint2*>(base);
    int8_t* eA = reinterpret_cast<int8_t*>(base + (((size_t)nA * nseg * kColSplit * sizeof(uint2) + 1023) & ~(size_t)1023));
    int8_t* eB = eA + pa * kRowBytes;
    CUtensorMap mapA, mapB;
    if (encode_desc_map(&mapA, eA, pa, kBM) || encode_desc_map(&mapB, eB, pb, kBN)) return -1;
    static size_t have[64] = {};
    if (opt_in_smem(match_tc_kernel, tc_smem_bytes(), have)) return -1;
    // rows past the end of either set: zeros (dot product 0); the epilogue never reads columns past nB, rows past nA are not written
    if (pa > (size_t)nA) cudaMemsetAsync(eA + (size_t)nA * kRowBytes, 0, (pa - nA) * kRowBytes, s);
    if (pb > (size_t)nB) cudaMemsetAsync(eB + (size_t)nB * kRowBytes, 0, (pb - nB) * kRowBytes, s);
    match_expand_kernel<<<(nA * 32 + 255) / 256, 256, 0, s>>>(dA, nA, eA);
    match_expand_kernel<<<(nB * 32 + 255) / 256, 256, 0, s>>>(dB, nB, eB);
    const int ntiles = (int)(pb / kBN), tiles_per_seg = (ntiles + nseg - 1) / nseg;
    dim3 grd((unsigned)(pa / kBM), (unsigned)nseg);
    match_tc_kernel<<<grd, kTcThreads, tc_smem_bytes(), s>>>(mapA, mapB, nA, nB, tiles_per_seg, part);
    launch_match_merge(part, nA, kColSplit * nseg, ratio, threshold, best_idx, best_dist, second_dist, match, s);
    return 0;
}

}  // namespace gforb
