// Integer-pipe issue-rate microbenchmark for sm_100a ("INT32/popc peak rates for B200 are not in
// MEASURED_PEAKS.json -> measure them").  Each kernel runs a dependent-chain-free unrolled loop of ONE instruction
// class on every SM at full occupancy and reports thread-ops per clock per SM and G thread-ops/s for the chip.
// The numbers are the INT/LSU roofline denominators quoted in DESIGN.md and bench.py ("int_roofline").
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/int_peaks tools/int_peaks.cu
// run  : tools/_build/int_peaks  > profiles/int_peaks_rN.json
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096;
constexpr int kIlp = 8;     // independent chains per thread

template <int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c) {
    if constexpr (OP == 0) return (a & b) ^ c;                                 // LOP3
    else if constexpr (OP == 1) return a + b + c;                              // IADD3
    else if constexpr (OP == 2) return __popc(a) + b;                          // POPC (+ IADD)
    else if constexpr (OP == 3) return __vabsdiffu4(a, b);                     // VABSDIFF4
    else if constexpr (OP == 4) return __vimax3_s16x2(a, b, c);                // VIMNMX3 (packed s16x2)
    else if constexpr (OP == 5) return __dp4a(a, b, c);                        // IDP4A
    else if constexpr (OP == 6) return a * b + c;                              // IMAD
    else if constexpr (OP == 7) return __funnelshift_r(a, b, c);               // SHF
    else if constexpr (OP == 8) return __byte_perm(a, b, c);                   // PRMT
    else if constexpr (OP == 9) return max(a, b) ^ c;                          // IMNMX + LOP3
    else return (a & b) * c + b;                                               // LOP3 + IMAD (ALU + FMA pipes together)
}

template <int OP> __global__ void __launch_bounds__(1024) rate_kernel(uint32_t* out, uint32_t seed) {
    uint32_t v[kIlp];
#pragma unroll
    for (int i = 0; i < kIlp; ++i) v[i] = seed * (threadIdx.x + i + 1);
    const uint32_t b = seed ^ 0x9E3779B9u, c = seed | 5u;
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int i = 0; i < kIlp; ++i) v[i] = op<OP>(v[i], OP == 2 ? b : v[(i + 1) % kIlp], c);   // cross-chain operand: nothing folds
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kIlp; ++i) s ^= v[i];
    if (s == 0x12345678u) out[0] = s;   // keep the chain alive
}

// shared-memory load rates: 32-bit (4 B/lane) and 128-bit (16 B/lane), conflict-free
template <int W> __global__ void __launch_bounds__(1024) lds_kernel(uint32_t* out, uint32_t seed) {
    __shared__ __align__(16) uint32_t sm[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = i * seed;
    __syncthreads();
    uint32_t acc = 0;
    int idx = threadIdx.x * W;
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int i = 0; i < kIlp; ++i) {
            const int j = (idx + i * 1024) & 8191 & ~(W - 1);
            if constexpr (W == 1) acc += sm[j];
            else { const uint4 q = *reinterpret_cast<const uint4*>(&sm[j]); acc += q.x ^ q.y ^ q.z ^ q.w; }
        }
        idx += acc & W;     // data-dependent so loads are not hoisted
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <typename F> static double time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount, ctas = sms * 2, thr = 1024;
    uint32_t* out; cudaMalloc(&out, 64);
    const char* names[11] = {"LOP3", "IADD3", "POPC+IADD", "VABSDIFF4", "VIMNMX3.S16x2", "IDP4A", "IMAD", "SHF", "PRMT", "IMNMX+LOP3 (2 ops)", "LOP3+IMAD (2 ops)"};
    double ms[13];
    ms[0] = time_ms([&] { rate_kernel<0><<<ctas, thr>>>(out, 3); });
    ms[1] = time_ms([&] { rate_kernel<1><<<ctas, thr>>>(out, 3); });
    ms[2] = time_ms([&] { rate_kernel<2><<<ctas, thr>>>(out, 3); });
    ms[3] = time_ms([&] { rate_kernel<3><<<ctas, thr>>>(out, 3); });
    ms[4] = time_ms([&] { rate_kernel<4><<<ctas, thr>>>(out, 3); });
    ms[5] = time_ms([&] { rate_kernel<5><<<ctas, thr>>>(out, 3); });
    ms[6] = time_ms([&] { rate_kernel<6><<<ctas, thr>>>(out, 3); });
    ms[7] = time_ms([&] { rate_kernel<7><<<ctas, thr>>>(out, 3); });
    ms[8] = time_ms([&] { rate_kernel<8><<<ctas, thr>>>(out, 3); });
    ms[9] = time_ms([&] { rate_kernel<9><<<ctas, thr>>>(out, 3); });
    ms[10] = time_ms([&] { rate_kernel<10><<<ctas, thr>>>(out, 3); });
    ms[11] = time_ms([&] { lds_kernel<1><<<ctas, thr>>>(out, 3); });
    ms[12] = time_ms([&] { lds_kernel<4><<<ctas, thr>>>(out, 3); });
    const double ops = (double)ctas * thr * kIters * kIlp;
    const double clk = clk_khz * 1e3;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_attr\": %.0f, \"note\": \"thread-ops; per_clk_per_sm assumes the max SM clock\",\n \"rates\": {", p.name, sms, clk / 1e6);
    for (int i = 0; i < 13; ++i) {
        const char* n = i < 11 ? names[i] : (i == 11 ? "LDS.32" : "LDS.128");
        const double gops = ops * (i == 9 || i == 10 ? 2 : 1) / (ms[i] * 1e-3) / 1e9;
        printf("%s\n  \"%s\": {\"ms\": %.4f, \"gops\": %.1f, \"per_clk_per_sm\": %.1f}", i ? "," : "", n, ms[i], gops,
               gops * 1e9 / clk / sms);
    }
    printf("\n }}\n");
    return 0;
}
