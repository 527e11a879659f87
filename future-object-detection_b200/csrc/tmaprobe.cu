// Probe: does a tiled tensor map whose second dimension's stride (16 B) is SMALLER than the first dimension's extent
// (32 elements = 64 B) encode and load as an overlapping-window gather?  That is the stride-2 7x7 stem convolution's
// im2col row for one kernel row when the image is stored NHWC with 4 channels (8 B per pixel): output pixel x reads
// the 8 pixels starting at input pixel 2x (left padding stored explicitly).  Rank 5: [k, x, row parity, row / 2, n].
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
#include "sm100_ptx.cuh"
#include "tmap.h"
using namespace fod;

__global__ void probe_kernel(const __grid_constant__ CUtensorMap tm, int x0, int par, int yy0, int n, uint16_t* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, 128 * 64);
        tma_load_5d(smem, &tm, &bar, 0, x0, par, yy0, n);
    }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < 128 * 32; i += blockDim.x) out[i] = reinterpret_cast<uint16_t*>(smem)[i];
}

int main() {
    const int N = 2, H = 16, WP = 40, OW = 16;  // (OW - 1) * 16 + 64 = 304 <= 320 = row pitch
    const int pitch = WP * 8;
    std::vector<uint16_t> h((size_t)N * H * WP * 4);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint16_t)(i + 1);
    uint16_t *d, *o;
    cudaMalloc(&d, h.size() * 2);
    cudaMalloc(&o, 128 * 64);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    CUtensorMap tm;
    uint64_t dims[5] = {32, (uint64_t)OW, 2, (uint64_t)H / 2, (uint64_t)N};
    uint64_t str[4] = {16, (uint64_t)pitch, 2ull * pitch, (uint64_t)H * pitch};
    uint32_t box[5] = {32, 16, 1, 8, 1};
    if (make_tmap_bf16(&tm, d, 5, dims, str, box, 64)) {
        printf("PROBE encode FAILED\n");
        return 1;
    }
    printf("PROBE encode ok\n");
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 64 + 1024);
    int bad_total = 0;
    const int cases[4][4] = {{0, 0, 0, 0}, {0, 1, -1, 1}, {0, 1, 2, 1}, {8, 0, 4, 0}};  // x0, par, yy0, n
    for (auto& c : cases) {
        cudaMemset(o, 0xff, 128 * 64);
        probe_kernel<<<1, 128, 128 * 64 + 1024>>>(tm, c[0], c[1], c[2], c[3], o);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("PROBE launch error: %s\n", cudaGetErrorString(e));
            return 2;
        }
        std::vector<uint16_t> r(128 * 32);
        cudaMemcpy(r.data(), o, 128 * 64, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int y = 0; y < 8; ++y)
            for (int x = 0; x < 16; ++x)
                for (int k = 0; k < 32; ++k) {
                    const int row = y * 16 + x;
                    const int chunk = k >> 3, phys = chunk ^ ((row >> 1) & 3);
                    const uint16_t got = r[row * 32 + phys * 8 + (k & 7)];
                    const int yy = c[2] + y, xo = c[0] + x;
                    uint16_t want = 0;
                    if (yy >= 0 && yy < H / 2 && xo < OW) {
                        const int irow = 2 * yy + c[1];
                        want = h[((size_t)c[3] * H + irow) * WP * 4 + xo * 8 + k];
                    }
                    if (got != want && bad++ < 4) printf("  mismatch y %d x %d k %d: got %u want %u\n", y, x, k, got, want);
                }
        printf("PROBE case x0 %d par %d yy0 %d n %d: %d mismatches\n", c[0], c[1], c[2], c[3], bad);
        bad_total += bad;
    }
    printf(bad_total ? "PROBE RESULT: FAIL\n" : "PROBE RESULT: overlapping-stride gather works\n");
    return 0;
}
