// gforb_match_bruteforce_static: the brute-force Hamming matcher of k_match.cu for callers that may not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// gforb_match_bruteforce_device keeps its per-segment (best, second) key pairs in a scratch buffer that the matcher
// handle cudaMallocs and grows on demand.  Here the same two kernels (match_scan_kernel -> match_merge_kernel, launched
// through the unchanged launch_match of k_match.cu: identical outputs) get their scratch from a file-scope device array
// of fixed size instead:
//   * kStaticPairs = 524,288 key pairs = 4 MiB, held by every GPU that loads the library until the process exits;
//   * n_a <= GFORB_MATCH_STATIC_MAX_QUERIES (65,536); n_b <= 4,194,304 as everywhere; more is refused with
//     GFORB_ERR_CAPACITY.  The train set is cut into min(what match_segments picks, kStaticPairs / n_a) segments, so
//     large query sets run with fewer, longer segments than the handle-based entry point would use (>= 8 at 65,536);
//   * the call makes no cudaMalloc / cudaFree / cudaStreamCreate / cudaEventCreate and takes no handle: inputs, outputs
//     and the stream are the caller's.
// In host code `g_match_scratch` is only a host-side stand-in for the device array: its device address is taken with
// cudaGetSymbolAddress, once per device and with that device current, and that pointer is what the kernels receive.
// Every call on a device shares the array.  It keeps the last call's contents, which is harmless: the scan kernel
// writes every pair the merge kernel of the same call reads (all n_a x nseg of them), nothing has to start at zero.
// Two calls whose kernels overlap on the GPU (different streams, no dependency between them) RACE on it: serialise them.
#include <mutex>

#include "engine.hpp"
#include "gforb_internal.cuh"

namespace gforb {
namespace {
constexpr int kStaticMaxQueries = GFORB_MATCH_STATIC_MAX_QUERIES;
constexpr size_t kStaticPairs = (size_t)8 * kStaticMaxQueries;        // 524,288 x 8 bytes = 4 MiB
constexpr int kMaxDevices = 64;

static __device__ __align__(256) uint2 g_match_scratch[kStaticPairs];

struct DeviceSlot { uint2* scratch = nullptr; int sms = 0; };
DeviceSlot g_slots[kMaxDevices];
std::mutex g_slots_mutex;

// the device address of g_match_scratch on the CURRENT device (looked up once per device) and that device's SM count
int static_scratch(uint2** ptr, int* sms) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) { set_error("gforb_match_bruteforce_static: device ordinal out of range"); return GFORB_ERR_ARG; }
    std::lock_guard<std::mutex> lock(g_slots_mutex);
    DeviceSlot& slot = g_slots[dev];
    if (!slot.scratch) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_match_scratch);
        if (e != cudaSuccess || !p) return cuda_fail(e != cudaSuccess ? e : cudaErrorInvalidSymbol, "cudaGetSymbolAddress(g_match_scratch)");
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute(multiProcessorCount)");
        slot.sms = n;
        slot.scratch = static_cast<uint2*>(p);
    }
    *ptr = slot.scratch;
    *sms = slot.sms;
    return GFORB_OK;
}
}  // namespace
}  // namespace gforb

using namespace gforb;

extern "C" GFORB_API int gforb_match_bruteforce_static(const uint8_t* dA, int nA, const uint8_t* dB, int nB, float nnratio,
                                                       int threshold, int32_t* best_idx, int32_t* best_dist,
                                                       int32_t* second_dist, int32_t* match, void* cuda_stream) {
    if (nA < 0 || nB < 0 || (nA > 0 && !dA) || (nB > 0 && !dB) || !(nnratio > 0.f)) {
        set_error("gforb_match_bruteforce_static: bad arguments");
        return GFORB_ERR_ARG;
    }
    if (nA > kStaticMaxQueries) {
        set_error("gforb_match_bruteforce_static: n_a exceeds 65,536 (GFORB_MATCH_STATIC_MAX_QUERIES), the bound of the static scratch; "
                  "use gforb_match_bruteforce_device");
        return GFORB_ERR_CAPACITY;
    }
    if (nB > match_max_train()) { set_error("gforb_match_bruteforce_static: n_b exceeds 4,194,304"); return GFORB_ERR_CAPACITY; }
    if (((uintptr_t)dA | (uintptr_t)dB) & 15) {
        set_error("gforb_match_bruteforce_static: descriptor arrays must be 16-byte aligned");
        return GFORB_ERR_ARG;
    }
    if (nA == 0) return GFORB_OK;
    uint2* scratch = nullptr;
    int sms = 0;
    const int rc = static_scratch(&scratch, &sms);
    if (rc) return rc;
    int nseg = match_segments(nA, nB, sms);
    nseg = (int)std::min<size_t>((size_t)nseg, kStaticPairs / (size_t)nA);       // n_a <= 65,536: at least 8
    if ((size_t)nA * (size_t)nseg > kStaticPairs || nseg < 1) {                   // cannot happen; the array is never overrun
        set_error("gforb_match_bruteforce_static: internal segment bound");
        return GFORB_ERR_CAPACITY;
    }
    launch_match(dA, nA, dB, nB, nnratio, threshold, nseg, scratch, best_idx, best_dist, second_dist, match,
                 (cudaStream_t)cuda_stream);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "match launch (static scratch)");
    return GFORB_OK;
}
