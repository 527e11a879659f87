// Device micro-benchmarks behind the roofline statements of DESIGN.md (standalone, no Python):
//   1. TMEM read bandwidth per SM (tcgen05.ld.32x32b.x32) with 4 / 8 / 16 reader warps — the ceiling of every
//      epilogue and of the attention softmax, which pull fp32 accumulators into registers;
//   2. ex2.approx throughput per SM (MUFU) next to a degree-3 polynomial exp2 on the FMA pipe;
//   3. tcgen05.mma issue cost by N, A operand in shared or tensor memory;
//   4. TMA ingest of the GEMM main loop (L2 -> shared memory) with and without the MMAs that consume it.
// Build: make microbench ; run on a B200: ./microbench_fod
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "sm100_ptx.cuh"
#include "tmap.h"

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


// L2 -> shared-memory ingest of the GEMM main loop: per k-block NA boxes of 128 rows x 64 k (16 KB each, this CTA's rows
// of A) and, with B, one 256 x 64 box of W (32 KB, the same for every CTA) through a 3- or 4-stage ring.  MMAS = 0: the
// consumer frees the stage at once (pure ingest); MMAS = m: it issues m M128 N256 K16 MMAs per 16 k first (operand reads
// compete with the TMA writes for shared-memory bandwidth).  Reports clk per k-block.
template <int NA, bool WITHB, int MMAS, int STAGES>
__global__ void __launch_bounds__(128, 1) ingest_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                         int reps, int kblocks, unsigned long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int kStage = NA * 16384 + (WITHB ? 32768 : 0);
    __shared__ uint64_t full[STAGES], empty[STAGES];
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full + i, 1);
            mbar_init(empty + i, 1);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<512>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot;
    const int m0 = blockIdx.x * 128 * NA;
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        int st = 0;
        uint32_t ph = 0;
        for (int r = 0; r < reps; ++r)
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(empty + st, ph ^ 1);
                uint8_t* sa = smem + st * kStage;
                mbar_arrive_expect_tx(full + st, kStage);
#pragma unroll
                for (int t = 0; t < NA; ++t) tma_load_2d(sa + t * 16384, &tmA, full + st, kb * 64, m0 + t * 128);
                if (WITHB) tma_load_2d(sa + NA * 16384, &tmB, full + st, kb * 64, 0);
                if (++st == STAGES) {
                    st = 0;
                    ph ^= 1;
                }
            }
    } else if (threadIdx.x == 32) {
        constexpr uint32_t idesc = umma_idesc_bf16(128, 256, 0, 0);
        int st = 0;
        uint32_t ph = 0;
        for (int r = 0; r < reps; ++r)
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(full + st, ph);
                if (MMAS > 0) {
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(smem) + st * kStage;
                    const uint64_t dA = umma_smem_desc(a0, 16, 1024, 2);
                    const uint64_t dB = umma_smem_desc(a0 + (WITHB ? NA * 16384 : 0), 16, 1024, 2);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int t = 0; t < MMAS; ++t)
                            umma_bf16_ss(base + (t & 1) * 256, umma_desc_adv(dA, (NA > 1 ? (t & 1) * 16384 : 0) + k * 32), umma_desc_adv(dB, k * 32), idesc, 1u);
                    }
                    umma_commit(empty + st);
                } else {
                    mbar_arrive(empty + st);
                }
                if (++st == STAGES) {
                    st = 0;
                    ph ^= 1;
                }
            }
        if (MMAS > 0) {  // drain: the last commit covers every MMA issued
            const int last = (st + STAGES - 1) % STAGES;
            const long long n = (long long)reps * kblocks;
            mbar_wait(empty + last, (uint32_t)(((n - 1) / STAGES) & 1));
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
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
    {
        printf("TMA ingest of the GEMM main loop (A rows per CTA, W shared), 32 k-blocks x 8 passes per CTA\n");
        const int rows = 148 * 256, K = 2048;
        void *dA, *dW;
        cudaMalloc(&dA, (size_t)rows * K * 2);
        cudaMalloc(&dW, (size_t)256 * K * 2);
        cudaMemset(dA, 0, (size_t)rows * K * 2);
        cudaMemset(dW, 0, (size_t)256 * K * 2);
        CUtensorMap tmA, tmB;
        uint64_t da[2] = {(uint64_t)K, (uint64_t)rows}, db[2] = {(uint64_t)K, 256};
        uint64_t str[1] = {(uint64_t)K * 2};
        uint32_t ba[2] = {64, 128}, bb[2] = {64, 256};
        if (fod::make_tmap_bf16(&tmA, dA, 2, da, str, ba, 128) || fod::make_tmap_bf16(&tmB, dW, 2, db, str, bb, 128)) return 1;
        const int reps = 8, kbs = 32;
#define IG(NA, WB, MM, ST, GRID)                                                                                      \
    {                                                                                                                  \
        const int smem = 1024 + ST * (NA * 16384 + (WB ? 32768 : 0));                                                  \
        cudaFuncSetAttribute(ingest_kernel<NA, WB, MM, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);        \
        cudaMemset(d_cyc, 0, 148 * 8);                                                                                 \
        double c = run([&] { ingest_kernel<NA, WB, MM, ST><<<GRID, 128, smem>>>(tmA, tmB, reps, kbs, d_cyc); }, d_cyc) * 148.0 / GRID; \
        const double per = c / (reps * kbs);                                                                           \
        printf("  grid %3d  A boxes %d  W %d  MMAs/k16 %d  stages %d: %.0f clk per k-block -> %.1f B/clk/SM, %.0f B/clk chip\n", GRID, NA, \
               (int)WB, MM, ST, per, (NA * 16384 + (WB ? 32768 : 0)) / per, GRID * (NA * 16384 + (WB ? 32768 : 0)) / per); \
    }
        IG(1, true, 0, 3, 148) IG(1, true, 0, 4, 148) IG(1, true, 0, 3, 88) IG(1, true, 0, 3, 27) IG(1, true, 0, 3, 1)
        IG(2, true, 0, 3, 88) IG(2, true, 0, 3, 148) IG(1, false, 0, 4, 148) IG(2, false, 0, 4, 148) IG(2, false, 0, 4, 27)
        IG(1, true, 1, 3, 148) IG(1, true, 1, 4, 148) IG(1, true, 1, 3, 27) IG(2, true, 2, 3, 88) IG(2, true, 2, 3, 27)
    }
    return 0;
}
