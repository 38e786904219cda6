// Multi-index hashing of 256-bit descriptors: the descriptor side of HASHING::MultiIndexHashing
// (reference include/Hashing.h:171-216, src/Hashing.cc) -- substring keys (hash :53-103), batched bucket insertion
// (insert :533-598 over Bucket::add :120-154) and the two candidate queries (:699-846, :849-950 over Bucket::back :250-279).
//
// Device layout (per table t of T): the reference's per-bucket std::deque with push_front is a singly linked chain,
// newest entry first --
//   head[t][2^bits]   slot of the bucket's newest entry (-1: empty); also Bucket::front()'s mLatest
//   pool_id[t][cap]   entry handle (MapPoint::mnId) held by a slot
//   pool_next[t][cap] slot of the next older entry of the same bucket (-1: end)
//   fill[t]           slots in use
// so Bucket::back(n) = the first n steps down the chain, and growing a table never moves an entry.
//
// insert: the reference inserts feature by feature; what its result depends on is only the ORDER of the entries inside
// each bucket and the "same id as the bucket's latest" test (:587).  One CTA per table sorts the batch's (key, batch
// position) pairs in shared memory (bitonic, <= 8,192 per launch), then one thread per bucket walks its run in batch order
// and links it in front of the old chain -- identical chains, no atomics on the chain itself.
// query: one warp per query descriptor, lane = table: chain walks are independent pointer chases (latency-bound; many
// queries in flight hide it), a warp scan places each table's run in the output in table order.
#include "gforb_internal.cuh"

namespace gforb {

namespace {
constexpr int kInsertThreads = 256;
constexpr int kPosBitsMih = 13;                    // batch positions 0..8191 below the key in the sort word

// MultiIndexHashing::hash (:53-103): table i = bits [i * bits, (i + 1) * bits) of the descriptor's 32-bit words, LSB first
__device__ __forceinline__ uint32_t mih_key(const uint32_t* __restrict__ d, int t, int bits) {
    const int b0 = bits * t, si = b0 >> 5, sp = b0 & 31, ei = (b0 + bits) >> 5, ep = (b0 + bits) & 31;
    const uint32_t mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
    if (si == ei || ep == 0) return (d[si] >> sp) & mask;                 // inside one word (or ending on its last bit)
    return (d[si] >> sp) | ((d[ei] & ((1u << ep) - 1u)) << (bits - ep));  // across two words
}
}  // namespace

__global__ void __launch_bounds__(256)
mih_hash_kernel(const uint32_t* __restrict__ desc, int n, int T, int bits, uint32_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * T) return;
    const int e = i / T, t = i - e * T;
    keys[i] = mih_key(desc + 8 * (size_t)e, t, bits);
}

__global__ void __launch_bounds__(kInsertThreads)
mih_insert_kernel(const uint32_t* __restrict__ desc, const int32_t* __restrict__ entry, int n, int n2, int bits, uint32_t B, int cap,
                  int32_t* __restrict__ head, int32_t* __restrict__ pool_id, int32_t* __restrict__ pool_next,
                  int32_t* __restrict__ fill, uint32_t* __restrict__ hashed, uint8_t* __restrict__ success, int32_t* __restrict__ err) {
    extern __shared__ uint32_t comp[];              // (key << 13 | batch position), 0xFFFFFFFF = not inserted into this table
    const int t = blockIdx.x;
    head += (size_t)t * B; pool_id += (size_t)t * cap; pool_next += (size_t)t * cap;
    for (int e = threadIdx.x; e < n2; e += kInsertThreads) {
        uint32_t c = 0xFFFFFFFFu;
        if (e < n) {
            const uint32_t old = atomicOr(&hashed[e], 1u << t);            // feature->mvbHashed[hashtable_id] (:571-573)
            if (!(old >> t & 1u)) c = (mih_key(desc + 8 * (size_t)e, t, bits) << kPosBitsMih) | (uint32_t)e;
        }
        comp[e] = c;
    }
    __syncthreads();
    for (int k = 2; k <= n2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n2; i += kInsertThreads) {
                const int p = i ^ j;
                if (p > i) {
                    const uint32_t a = comp[i], b = comp[p];
                    if ((a > b) == ((i & k) == 0)) { comp[i] = b; comp[p] = a; }
                }
            }
            __syncthreads();
        }
    // one thread per bucket that received entries: link its run, in batch order, in front of the old chain
    for (int i = threadIdx.x; i < n2; i += kInsertThreads) {
        const uint32_t c = comp[i];
        if (c == 0xFFFFFFFFu) continue;
        const uint32_t key = c >> kPosBitsMih;
        if (i > 0 && (comp[i - 1] >> kPosBitsMih) == key) continue;
        int latest = head[key];
        int latest_id = latest >= 0 ? pool_id[latest] : 0;
        for (int j = i; j < n2; ++j) {
            const uint32_t cj = comp[j];
            if (cj == 0xFFFFFFFFu || (cj >> kPosBitsMih) != key) break;
            const int e = (int)(cj & ((1u << kPosBitsMih) - 1u));
            const int id = entry[e];
            if (latest < 0 || latest_id != id) {                           // latest == NULL || latest->mnId != feature->mnId (:587)
                const int slot = atomicAdd(fill + t, 1);
                if (slot >= cap) { atomicOr(err, 1); break; }              // cannot happen: the host reserves n slots per table
                pool_id[slot] = id;
                pool_next[slot] = latest;
                latest = slot; latest_id = id;
                success[e] = 1;
            }
        }
        head[key] = latest;
    }
}

struct MihQueryArgs {
    const uint32_t* qdesc; int m, T, bits; uint32_t B; int cap;
    const int32_t* head; const int32_t* pool_id; const int32_t* pool_next;
    int take[32];                 // Bucket::back's n for table k, already >= 0 (0: table not consulted)
    int keep;                     // NUM_CANDIDATE_FEATURES + dynamics for the split form; INT_MAX for the plain form
    int32_t* cand; int32_t* count; int32_t* addi; int32_t* addi_count; int stride;
};

__global__ void __launch_bounds__(256)
mih_query_kernel(const MihQueryArgs a) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= a.m) return;
    const bool on = lane < a.T && a.take[lane] > 0;
    int first = -1, c = 0;
    const int32_t* next = a.pool_next + (size_t)lane * a.cap;
    const int32_t* ids = a.pool_id + (size_t)lane * a.cap;
    if (on) {
        first = a.head[(size_t)lane * a.B + mih_key(a.qdesc + 8 * (size_t)q, lane, a.bits)];
        for (int s = first; s >= 0 && c < a.take[lane]; s = next[s]) ++c;
    }
    // the newest c - keep entries go to the additional list, the oldest `keep` of the run to the candidates (:884-893)
    const int nadd = c > a.keep ? c - a.keep : 0, ncand = c - nadd;
    int ic = ncand, ia = nadd;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int tc = __shfl_up_sync(0xffffffffu, ic, o), ta = __shfl_up_sync(0xffffffffu, ia, o);
        if (lane >= o) { ic += tc; ia += ta; }
    }
    if (lane == 31) {
        a.count[q] = ic;
        if (a.addi_count) a.addi_count[q] = ia;
    }
    int32_t* oc = a.cand + (size_t)q * a.stride + (ic - ncand);
    int32_t* oa = a.addi ? a.addi + (size_t)q * a.stride + (ia - nadd) : nullptr;
    int j = 0;
    for (int s = first; j < c; s = next[s], ++j) {
        if (j < nadd) oa[j] = ids[s];
        else oc[j - nadd] = ids[s];
    }
}

void launch_mih_hash(const uint8_t* desc, int n, int T, int bits, uint32_t* keys, cudaStream_t s) {
    if (n > 0) mih_hash_kernel<<<(n * T + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(desc), n, T, bits, keys);
}

int mih_max_batch() { return 1 << kPosBitsMih; }

// n <= mih_max_batch(); hashed / success: n entries; returns -1 when the shared-memory opt-in fails
int launch_mih_insert(const uint8_t* desc, const int32_t* entry, int n, int T, int bits, int cap, int32_t* head, int32_t* pool_id,
                      int32_t* pool_next, int32_t* fill, uint32_t* hashed, uint8_t* success, int32_t* err, cudaStream_t s) {
    if (n <= 0) return 0;
    int n2 = 32;
    while (n2 < n) n2 <<= 1;
    static size_t have[64] = {};
    const size_t smem = (size_t)n2 * 4;
    if (smem > 32 * 1024 && opt_in_smem(mih_insert_kernel, smem, have)) return -1;       // static + dynamic share the 48 KB default
    mih_insert_kernel<<<T, kInsertThreads, smem, s>>>(reinterpret_cast<const uint32_t*>(desc), entry, n, n2, bits, 1u << bits, cap, head,
                                                      pool_id, pool_next, fill, hashed, success, err);
    return 0;
}

void launch_mih_query(const uint8_t* qdesc, int m, int T, int bits, int cap, const int32_t* head, const int32_t* pool_id,
                      const int32_t* pool_next, const int* take, int keep, int32_t* cand, int32_t* count, int32_t* addi,
                      int32_t* addi_count, int stride, cudaStream_t s) {
    if (m <= 0) return;
    MihQueryArgs a{};
    a.qdesc = reinterpret_cast<const uint32_t*>(qdesc); a.m = m; a.T = T; a.bits = bits; a.B = 1u << bits; a.cap = cap;
    a.head = head; a.pool_id = pool_id; a.pool_next = pool_next;
    for (int k = 0; k < 32; ++k) a.take[k] = k < T ? take[k] : 0;
    a.keep = keep; a.cand = cand; a.count = count; a.addi = addi; a.addi_count = addi_count; a.stride = stride;
    mih_query_kernel<<<(m + 7) / 8, 256, 0, s>>>(a);
}

}  // namespace gforb
