// TMA (cp.async.bulk.tensor) + mbarrier helpers for sm_100a: image tiles global -> shared.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace gforb {

// Host: 3-D u8 tensor map {x = pitch bytes, y = padded rows, z = images} with a (box_w x box_h x 1) box.
// Returns 0 on success, else a CUresult.  Resolved through cudaGetDriverEntryPoint (no libcuda link).
int encode_tile_map_u8(CUtensorMap* map, const void* base, uint64_t pitch, uint64_t rows, uint64_t nimg,
                       uint32_t box_w, uint32_t box_h);

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
// one (box_w x box_h) u8 tile at element coordinates (x, y, z) -> dst (128-byte aligned shared memory)
__device__ __forceinline__ void tma_load_tile_3d(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)) : "memory");
}
#endif

}  // namespace gforb
