// Device micro-benchmarks behind the roofline statements of DESIGN.md (standalone, no Python):
//   1. TMEM read bandwidth per SM (tcgen05.ld.32x32b.x32) with 4 / 8 / 16 reader warps — the ceiling of every
//      epilogue and of the attention softmax, which pull fp32 accumulators into registers;
//   2. ex2.approx throughput per SM (MUFU) next to a degree-3 polynomial exp2 on the FMA pipe.
// Build: make microbench ; run on a B200: ./microbench_fod
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "sm100_ptx.cuh"

using namespace fod;

template <int kWarps, int kBatch>
__global__ void __launch_bounds__(kWarps * 32, 1) tmem_read_kernel(int iters, unsigned long long* cycles, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc<512>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot;
    const uint32_t taddr = base + (uint32_t((warp & 3) * 32) << 16);
    uint32_t v[32];
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int b = 0; b < kBatch; ++b) {
            tmem_ld_32x32b_x32(taddr + ((i * kBatch + b) & 15) * 32, v);
            if (b == kBatch - 1) tmem_ld_wait();
            acc ^= v[b & 31];
        }
    }
    tmem_ld_wait();
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0x12345678u) sink[0] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<512>(base);
    }
}

template <int kMode>
__global__ void __launch_bounds__(512, 1) exp_kernel(int iters, unsigned long long* cycles, float* sink) {
    float x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = -0.001f * (threadIdx.x + j);
    float accv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float y;
            if (kMode == 0) {
                y = fast_exp2(x[j]);
            } else {
                // Cody-Waite: 2^x = 2^floor(x) * p(frac), p = degree-3 minimax on [0,1)
                const float fl = floorf(x[j]);
                const float f = x[j] - fl;
                float p = fmaf(f, 0.0790210605f, 0.2240446399f);
                p = fmaf(p, f, 0.6967699046f);
                p = fmaf(p, f, 1.0f);
                y = __int_as_float(__float_as_int(p) + (int(fl) << 23));
            }
            accv[j] += y;
            x[j] -= 1e-4f;
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += accv[j];
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 123.456f) sink[0] = acc;
}

// Issue rate of tcgen05.mma (M = 128, K = 16, bf16) by N, with the A operand in shared memory (SS) or in tensor memory (TS):
// one thread issues kPer MMAs, commits, waits; smem / TMEM contents are whatever is there (timing only).
template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int iters, unsigned long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<512>(&slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot;
    constexpr int kPer = 32;
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
        const uint64_t dA = umma_smem_desc(smem_u32(smem), 16, 1024, 2);            // 128 x 64 bf16, K-major, SWIZZLE_128B
        const uint64_t dB = umma_smem_desc(smem_u32(smem) + 16384, 16, 1024, 2);    // N x 64
        uint32_t ph = 0;
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < kPer; ++k) {
                if (TS) umma_bf16_ts(base, base + 256 + (k & 3) * 8, umma_desc_adv(dB, (k & 3) * 32), idesc, 1u);
                else umma_bf16_ss(base, umma_desc_adv(dA, (k & 3) * 32), umma_desc_adv(dB, (k & 3) * 32), idesc, 1u);
            }
            umma_commit(&bar);
            mbar_wait(&bar, ph);
            ph ^= 1;
        }
        const long long t1 = clock64();
        cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc<512>(base);
    }
}

template <typename F>
static double run(F launch, unsigned long long* d_cyc) {
    launch();
    cudaDeviceSynchronize();
    launch();
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        printf("CUDA error: %s\n", cudaGetErrorString(e));
        return -1;
    }
    unsigned long long h[148];
    cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double s = 0;
    for (int i = 0; i < 148; ++i) s += (double)h[i];
    return s / 148;
}

int main() {
    unsigned long long* d_cyc;
    uint32_t* d_sink;
    cudaMalloc(&d_cyc, 148 * 8);
    cudaMalloc(&d_sink, 64);
    const int iters = 2000;
    printf("TMEM read (tcgen05.ld.32x32b.x32 = 4 KB per warp instruction), 148 CTAs, %d iterations\n", iters);
#define TM(W, B)                                                                                             \
    {                                                                                                        \
        double c = run([&] { tmem_read_kernel<W, B><<<148, W * 32>>>(iters, d_cyc, d_sink); }, d_cyc);       \
        printf("  warps %2d, %d loads per wait: %.0f cycles -> %.1f B/clk/SM, %.1f clk per warp-load\n", W, B, c, \
               (double)iters * B * W * 4096.0 / c, c / ((double)iters * B));                                 \
    }
    TM(4, 1) TM(4, 4) TM(8, 1) TM(8, 4) TM(16, 1) TM(16, 4)
    printf("exp2 throughput, 148 CTAs x 512 threads\n");
    float* fs = reinterpret_cast<float*>(d_sink);
    {
        double c = run([&] { exp_kernel<0><<<148, 512>>>(iters, d_cyc, fs); }, d_cyc);
        printf("  MUFU ex2.approx : %.0f cycles -> %.2f exp/clk/SM\n", c, (double)iters * 8 * 512 / c);
        c = run([&] { exp_kernel<1><<<148, 512>>>(iters, d_cyc, fs); }, d_cyc);
        printf("  FMA-pipe poly-3 : %.0f cycles -> %.2f exp/clk/SM\n", c, (double)iters * 8 * 512 / c);
    }
    printf("tcgen05.mma issue rate, M128 K16 bf16, 148 CTAs, 32 MMAs per commit\n");
    const int mi = 200;
#define MR(NN, TSV)                                                                                             \
    {                                                                                                           \
        cudaFuncSetAttribute(mma_rate_kernel<NN, TSV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 52000);     \
        double c = run([&] { mma_rate_kernel<NN, TSV><<<148, 128, 52000>>>(mi, d_cyc); }, d_cyc);               \
        printf("  N %3d  A in %s: %.1f clk per MMA -> %.0f flop/clk/SM\n", NN, TSV ? "TMEM" : "smem", c / (mi * 32.0), \
               2.0 * 128 * NN * 16 / (c / (mi * 32.0)));                                                        \
    }
    MR(32, false) MR(64, false) MR(128, false) MR(256, false) MR(32, true) MR(64, true) MR(128, true) MR(256, true)
    return 0;
}
