// Stand-alone probe of the TMA tile-load helper (gf-orb-slam2_b200/csrc/tma.cuh): loads one (box_w x box_h)
// u8 tile of a {pitch, rows, images} tensor at an arbitrary coordinate and compares with the source.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tma_probe tests/cpp/tma_probe.cu gf-orb-slam2_b200/csrc/_build/engine.o ... (see run line in DESIGN.md)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../gf-orb-slam2_b200/csrc/tma.cuh"

using namespace gforb;
struct alignas(64) Maps { CUtensorMap m[3]; };

__global__ void probe(const __grid_constant__ Maps maps, int which, int x, int y, int z, int bw, int bh, uint8_t* out) {
    __shared__ __align__(128) uint8_t tile[80 * 66];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) mbar_init(&bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, bw * bh);
        tma_load_tile_3d(tile, &maps.m[which], x, y, z, &bar);
    }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = tile[i];
}

int main() {
    const int pitch = 832, rows = 518, nimg = 3, bw = 48, bh = 38;
    std::vector<uint8_t> h((size_t)pitch * rows * nimg);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 24);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, bw * bh);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    Maps maps;
    for (int i = 0; i < 3; ++i) {
        int rc = encode_tile_map_u8(&maps.m[i], d, pitch, rows, nimg, bw, bh);
        if (rc) { printf("encode failed %d\n", rc); return 1; }
    }
    int bad_total = 0;
    // the innermost coordinate must be a multiple of 16 bytes: (77,35,0) raises cudaErrorIllegalInstruction on B200
    const int coords[5][3] = {{48, 35, 0}, {64, 101, 1}, {800, 500, 2}, {0, 0, 2}, {816, 517, 1}};
    for (auto& c : coords) {
        probe<<<1, 128>>>(maps, 1, c[0], c[1], c[2], bw, bh, o);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("tile (%d,%d,%d): kernel error: %s\n", c[0], c[1], c[2], cudaGetErrorString(e)); return 2; }
        std::vector<uint8_t> r(bw * bh);
        cudaMemcpy(r.data(), o, r.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int yy = 0; yy < bh; ++yy) for (int xx = 0; xx < bw; ++xx) {
            const int gx = c[0] + xx, gy = c[1] + yy;
            const uint8_t want = (gx < pitch && gy < rows) ? h[((size_t)c[2] * rows + gy) * pitch + gx] : 0;
            bad += r[yy * bw + xx] != want;
        }
        printf("tile (%d,%d,%d): %d mismatches\n", c[0], c[1], c[2], bad);
        bad_total += bad;
    }
    printf(bad_total ? "TMA PROBE FAILED\n" : "TMA PROBE OK\n");
    return bad_total != 0;
}
