// Y = epilogue(X · Wᵀ): warp-specialised TMA -> tcgen05.mma -> TMEM -> fused epilogue GEMM for sm_100a.
//   warp 0: TMA producer (A = X tile 128x64, B = W tile BNx64, SWIZZLE_128B, K-major both)
//   warp 1: TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=BN, K=16 per instruction)
//   warps 2..5: epilogue, one accumulator row per thread (tcgen05.ld 32x32b), bias / residual / ReLU /
//               LayerNorm over the full 256-wide row without leaving the CTA.
// Roofline: tensor-bound for the encoder FFN (K=256/2048, M>=5600); weight/launch bound for the decoder.
#include <cstdlib>
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"
#include "dropout.cuh"

namespace fod {

struct GemmEpi {
    const float* bias;
    const __nv_bfloat16* aux;        // residual (aux_mode 1) or ReLU mask (aux_mode 2) tile source, prefetched via cp.async
    const float* col_scale;          // optional (aux_row_mod or M, N) fp32 multiplier applied after bias: y = (acc+b)*s
    const float* ln_gamma;
    const float* ln_beta;
    __nv_bfloat16* out_bf16;
    float* out_f32;
    __nv_bfloat16* pre_ln;
    float* ln_mean;
    float* ln_rstd;
    float* colsum_out;               // optional (N) fp32: += column sums of the stored bf16 output (bias gradients)
    long long ld_aux, ld_out;
    int M, N, K;
    int aux_mode;                    // 0 none, 1 add residual, 2 zero where mask <= 0
    int aux_row_mod;                 // residual row = row % aux_row_mod when > 0
    int relu;
    int atomic_out;                  // 1: red.add into out_f32 (split-K / gradient accumulation)
    int kb_per_split;
    int num_stages;
    int vec_ok;                      // 1: 16-byte vector paths are legal for out / aux (alignment + N % 8 == 0)
    float ln_eps;
    float mask_scale;                // aux_mode 2: kept elements are multiplied by this (1/(1-p) of a dropped-out ReLU output)
    DropoutCfg drop;                 // state != nullptr: dropout on (acc + bias [ReLU]) — before the residual of the LN epilogue
};

// dropout on 32 consecutive columns [col0, col0 + 32) of global row `row` (element index row * N + col, N % 8 == 0)
__device__ __forceinline__ void drop_chunk(float (&z)[32], const PhiloxKey& key, const DropoutCfg& d, long long row, int N, int col0) {
    const unsigned long long g0 = ((unsigned long long)row * (unsigned long long)N + (unsigned long long)col0) >> 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        dropout_apply8(&z[q * 8], key, g0 + q, d.thresh, d.scale);
    }
}

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS_MAX = 320;
// Epilogue warps: 4 (one per TMEM lane quarter) for the 64/128-wide tiles; 8 for the 256-wide tile, where two warps
// share a lane quarter and split the columns — the one-row-per-thread epilogue is the longest phase of that tile.
template <int BN> struct GemmThreads { static constexpr int kEpiWarps = BN == 256 ? 8 : 4; static constexpr int kThreads = 64 + 32 * kEpiWarps; };
constexpr int GEMM_MAX_STAGES = 8;

// MT = 2: one CTA computes TWO 128-row tiles against the same BN x 64 k-blocks of W (two accumulators in TMEM, 512
// columns).  The N = 256 GEMMs over all encoder tokens (22 400 rows = 175 tiles on 148 SMs) take two waves, so the
// busiest SMs issue two tiles' MMAs whatever the tiling — the main loop itself runs at the MMA issue rate (csrc/
// microbench.cu, ingest_kernel: 541 clk per 48 KB k-block with the MMAs, 417 clk without; W, shared by every CTA, is
// served far above the ~7 KB/clk the L2 gives distinct lines).  What the two-tile CTA saves is the second CTA's launch,
// barrier / TMEM set-up and first TMA round trip, and half of the W fetches: -3..5 us on the 40..53 us K = 2048 calls.
template <int BN, int MT = 1>
struct GemmCfg {
    static constexpr int kABytes = MT * GEMM_BM * GEMM_BK * 2;
    static constexpr int kBBytes = BN * GEMM_BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    // epilogue staging rows carry 16 bytes of padding: a lane's 16-byte accesses to ITS row and a warp's accesses
    // along ONE row are both bank-conflict free (row stride = 4 banks mod 32)
    static constexpr int kRowBf16 = BN * 2 + 16;
    static constexpr int kRowF32 = BN * 4 + 16;
    static constexpr int kAuxBytes = GEMM_BM * kRowBf16;
    static constexpr int kTmemCols = BN < 32 ? 32 : MT * BN;
    static constexpr int kParamBytes = 3 * BN * 4;  // bias / gamma / beta slices of this N tile, fp32
    static constexpr int kStatBytes = MT * 2 * 2 * GEMM_BM * 4;  // LayerNorm partial sums of the two column halves, per sub-tile
    static constexpr int kBarBytes = 256 + kParamBytes + kStatBytes;
    // MT = 2 runs exactly 3 stages (3 x 64 KB) and has no room for a dedicated residual / mask staging area: once the
    // pipeline memory dies it is carved into [output staging | aux of sub-tile 1 | aux of sub-tile 0], each kAuxBytes.
    // The aux-0 region lies inside stage 2 (and a few KB past the pipeline), so its prefetch can start as soon as the
    // LAST use of stage 2 has retired — two or three k-blocks before the main loop ends.
    static constexpr int kMt2Stages = 3;
    static constexpr int kMt2Carve = 3 * kAuxBytes > kMt2Stages * kStageBytes ? 3 * kAuxBytes : kMt2Stages * kStageBytes;
    static int smem_bytes(int stages, bool aux) {
        if (MT == 2) return 1024 + kMt2Carve + kBarBytes;
        return 1024 + stages * kStageBytes + (aux ? kAuxBytes : 0) + kBarBytes;
    }
};

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}

// bias, col_scale, staged residual, ReLU, staged mask on 32 accumulator columns [col0, col0+32) of one row
template <int BN>
__device__ __forceinline__ void epi_chunk(float (&z)[32], const GemmEpi& p, const uint8_t* aux_row, long long row,
                                          long long mod_row, int col0, int c, const float* s_bias, const PhiloxKey& key) {
    if (p.bias) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float4 b4 = *reinterpret_cast<const float4*>(s_bias + c * 32 + q * 4);  // smem broadcast
            z[q * 4] += b4.x;
            z[q * 4 + 1] += b4.y;
            z[q * 4 + 2] += b4.z;
            z[q * 4 + 3] += b4.w;
        }
    }
    if (p.col_scale) {
        const float* sc = p.col_scale + mod_row * (long long)p.N + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (col0 + i < p.N) z[i] *= __ldg(sc + i);
    }
    if (p.aux_mode == 1) {
        if (p.vec_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(aux_row + c * 64 + q * 16), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[q * 8 + j] += f[j];
            }
        } else {
            const __nv_bfloat16* r = p.aux + mod_row * p.ld_aux + col0;
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) z[i] += __bfloat162float(r[i]);
        }
    }
    if (p.relu) {
#pragma unroll
        for (int i = 0; i < 32; ++i) z[i] = fmaxf(z[i], 0.f);
    }
    if (p.aux_mode == 2) {
        if (p.vec_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(aux_row + c * 64 + q * 16), f);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (!(f[j] > 0.f)) z[q * 8 + j] = 0.f;
            }
        } else {
            const __nv_bfloat16* mk = p.aux + row * p.ld_aux + col0;
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N && !(__bfloat162float(mk[i]) > 0.f)) z[i] = 0.f;
        }
        if (p.mask_scale != 1.f) {
#pragma unroll
            for (int i = 0; i < 32; ++i) z[i] *= p.mask_scale;
        }
    }
    if (p.drop.state) drop_chunk(z, key, p.drop, row, p.N, col0);  // (host: no residual together with dropout here)
}

__device__ __forceinline__ void stage_bf16(uint8_t* dst, const float (&z)[32]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(z[q * 8 + 0], z[q * 8 + 1]);
        u.y = pack_bf16x2(z[q * 8 + 2], z[q * 8 + 3]);
        u.z = pack_bf16x2(z[q * 8 + 4], z[q * 8 + 5]);
        u.w = pack_bf16x2(z[q * 8 + 6], z[q * 8 + 7]);
        reinterpret_cast<uint4*>(dst)[q] = u;
    }
}
__device__ __forceinline__ void stage_f32(uint8_t* dst, const float (&z)[32]) {
#pragma unroll
    for (int q = 0; q < 8; ++q)
        reinterpret_cast<float4*>(dst)[q] = make_float4(z[q * 4], z[q * 4 + 1], z[q * 4 + 2], z[q * 4 + 3]);
}

// One warp drains ITS 32 staged rows (bf16) to global memory with lane-contiguous 16-byte stores: a full 128-byte
// line per 8 lanes instead of 32 scattered row segments.  Optionally accumulates column sums of what it stored.
template <int BN, int W>
__device__ __forceinline__ void drain_bf16(const uint8_t* stage_warp, __nv_bfloat16* out, long long ld_out, int M, int N,
                                           long long row0, int n0, int lane, bool vec_ok, float* colsum_out) {
    // stage_warp points at this warp's first row AND first column; n0 is the global column of that first column
    constexpr int ROWB = GemmCfg<BN>::kRowBf16;
    if (vec_ok) {
        constexpr int CH = W / 8;     // 16-byte chunks per row: 32 / 16 / 8
        constexpr int RPI = 32 / CH;  // rows per warp-wide instruction
        const int ch = lane % CH, rsub = lane / CH;
        const int col = n0 + ch * 8;
        float cs[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) cs[j] = 0.f;
        if (col < N) {
#pragma unroll 4
            for (int r = rsub; r < 32; r += RPI) {
                if (row0 + r >= M) break;
                const uint4 u = *reinterpret_cast<const uint4*>(stage_warp + r * ROWB + ch * 16);
                *reinterpret_cast<uint4*>(out + (row0 + r) * ld_out + col) = u;
                if (colsum_out) {
                    float f[8];
                    unpack8(u, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) cs[j] += f[j];
                }
            }
        }
        if (colsum_out) {
#pragma unroll
            for (int o = CH; o < 32; o <<= 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], o);
            }
            if (rsub == 0 && col < N) {
#pragma unroll
                for (int j = 0; j < 8; ++j) atomicAdd(colsum_out + col + j, cs[j]);
            }
        }
    } else {
        for (int r = 0; r < 32; ++r) {
            if (row0 + r >= M) break;
            const __nv_bfloat16* s = reinterpret_cast<const __nv_bfloat16*>(stage_warp + r * ROWB);
            for (int cc = lane; cc < W; cc += 32) {
                if (n0 + cc < N) {
                    out[(row0 + r) * ld_out + n0 + cc] = s[cc];
                    if (colsum_out) atomicAdd(colsum_out + n0 + cc, __bfloat162float(s[cc]));
                }
            }
        }
    }
}

template <int BN, int W>
__device__ __forceinline__ void drain_f32(const uint8_t* stage_warp, float* out, long long ld_out, int M, int N,
                                          long long row0, int n0, int lane, bool vec_ok, bool atomic) {
    constexpr int ROWB = GemmCfg<BN>::kRowF32;
    if (vec_ok) {
        constexpr int CH = W / 4;  // float4 chunks per row
#pragma unroll 4
        for (int i = lane; i < 32 * CH; i += 32) {
            const int r = i / CH, ch = i % CH;
            if (row0 + r >= M) break;
            const int col = n0 + ch * 4;
            if (col >= N) continue;
            const float4 v = *reinterpret_cast<const float4*>(stage_warp + r * ROWB + ch * 16);
            float4* dst = reinterpret_cast<float4*>(out + (row0 + r) * ld_out + col);
            if (atomic) atomicAdd(dst, v);  // sm_90+: one 16-byte vector reduction
            else *dst = v;
        }
    } else {
        for (int r = 0; r < 32; ++r) {
            if (row0 + r >= M) break;
            const float* s = reinterpret_cast<const float*>(stage_warp + r * ROWB);
            for (int cc = lane; cc < W; cc += 32) {
                if (n0 + cc < N) {
                    if (atomic) atomicAdd(out + (row0 + r) * ld_out + n0 + cc, s[cc]);
                    else out[(row0 + r) * ld_out + n0 + cc] = s[cc];
                }
            }
        }
    }
}

// acc + bias + residual of 32 columns of one row for the LayerNorm epilogue (the only terms that epilogue admits)
__device__ __forceinline__ void ln_chunk(float (&z)[32], const uint32_t (&v)[32], const GemmEpi& p, const uint8_t* aux_row,
                                         long long row, long long mod_row, int c, const float* s_bias, bool row_ok,
                                         const PhiloxKey& key) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float4 b4 = *reinterpret_cast<const float4*>(s_bias + c * 32 + q * 4);  // zeros when there is no bias
        z[q * 4] = __uint_as_float(v[q * 4]) + b4.x;
        z[q * 4 + 1] = __uint_as_float(v[q * 4 + 1]) + b4.y;
        z[q * 4 + 2] = __uint_as_float(v[q * 4 + 2]) + b4.z;
        z[q * 4 + 3] = __uint_as_float(v[q * 4 + 3]) + b4.w;
    }
    if (p.drop.state && row_ok) drop_chunk(z, key, p.drop, row, p.N, c * 32);  // x + dropout(sublayer(x)): before the residual
    if (p.aux_mode == 1 && row_ok) {
        if (p.vec_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(aux_row + c * 64 + q * 16), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) z[q * 8 + j] += f[j];
            }
        } else {
            const __nv_bfloat16* r = p.aux + mod_row * p.ld_aux + c * 32;
#pragma unroll
            for (int i = 0; i < 32; ++i) z[i] += __bfloat162float(r[i]);
        }
    }
}

// LN = true: the LayerNorm epilogue (N == BN == 256) and nothing else — a separate instantiation, so that neither
// variant drags the other's code through the instruction cache.
template <int BN, bool A_MN, bool B_MN, bool LN, int MT = 1>
__global__ void __launch_bounds__(GemmThreads<BN>::kThreads, BN == 256 ? 1 : 2)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmEpi p) {
    using Cfg = GemmCfg<BN, MT>;
    static_assert(MT == 1 || (MT == 2 && BN == 256 && !A_MN), "two-tile CTAs: 256-wide tiles with a K-major A operand");
    static_assert(MT == 1 || 2 * Cfg::kAuxBytes >= 2 * Cfg::kStageBytes, "aux-0 region must start inside the last stage");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    // MT = 1: dedicated aux staging behind the stages.  MT = 2: aux of sub-tile t at (2 - t) * kAuxBytes of the carve.
    uint8_t* aux_stage = MT == 2 ? smem + 2 * Cfg::kAuxBytes : smem + p.num_stages * Cfg::kStageBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(MT == 2 ? smem + Cfg::kMt2Carve : aux_stage + (p.aux_mode ? Cfg::kAuxBytes : 0));
    uint64_t* empty_bar = full_bar + GEMM_MAX_STAGES;
    uint64_t* accum_bar = empty_bar + GEMM_MAX_STAGES;
    uint64_t* tail_bar = accum_bar + 1;  // MT = 2: the last MMAs that read stage 2 have retired
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail_bar + 1);
    float* s_bias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full_bar) + 256);
    float* s_gamma = s_bias + BN;
    float* s_beta = s_gamma + BN;
    float* s_stat = s_beta + BN;  // [MT sub-tiles][2 passes][2 halves][128 rows]
    constexpr int EW = GemmThreads<BN>::kEpiWarps;
    constexpr int CPW = (BN / 32) / (EW / 4);  // 32-column chunks per epilogue warp
    constexpr int WCOLS = CPW * 32;            // columns per epilogue warp

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const int m0 = blockIdx.y * (GEMM_BM * MT);
    const int total_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
    const int kb_begin = blockIdx.z * p.kb_per_split;
    const int kb_end = min(total_kb, kb_begin + p.kb_per_split);
    const int num_kb = kb_end - kb_begin;  // >= 1 by construction of the grid
    const int num_stages = p.num_stages;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < num_stages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        mbar_init(accum_bar, 1);
        mbar_init(tail_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor prefetch) may overlap
    // the tail of the previous kernel in the stream; nothing below reads or writes global memory before the previous
    // kernel has completed.  The next kernel's prologue may start as soon as every CTA of this grid is resident.
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(empty_bar + s, ph ^ 1);
                uint8_t* sa = smem + s * Cfg::kStageBytes;
                uint8_t* sb = sa + Cfg::kABytes;
                mbar_arrive_expect_tx(full_bar + s, Cfg::kStageBytes);
                const int k0 = (kb_begin + kb) * GEMM_BK;
                if constexpr (A_MN) {  // memory is (K, M): 64x64 boxes, one per 64-wide M chunk
#pragma unroll
                    for (int c = 0; c < GEMM_BM / 64; ++c) tma_load_2d(sa + c * 8192, &tmA, full_bar + s, m0 + c * 64, k0);
                } else {
#pragma unroll
                    for (int t = 0; t < MT; ++t)  // rows past M are zero-filled and still count their bytes
                        tma_load_2d(sa + t * (GEMM_BM * GEMM_BK * 2), &tmA, full_bar + s, k0, m0 + t * GEMM_BM);
                }
                if constexpr (B_MN) {
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, full_bar + s, n0 + c * 64, k0);
                } else {
                    tma_load_2d(sb, &tmB, full_bar + s, k0, n0);
                }
                if (++s == num_stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            // K-major: 128-byte rows of 64 k, 8-row groups 1024 B apart, 16 k per instruction = +32 B.  MN-major: [64 k][64 mn]
            // slabs of 8 KB (LBO = next 64 mn), 8-k groups 1024 B apart (SBO), 16 k per instruction = +2048 B.
            // Stage-0 descriptors are built once and stepped with one add per instruction (umma_desc_adv).
            const uint32_t a0 = smem_u32(smem);
            const uint64_t dA0 = A_MN ? umma_smem_desc(a0, 8192, 1024, 2) : umma_smem_desc(a0, 16, 1024, 2);
            const uint64_t dB0 = B_MN ? umma_smem_desc(a0 + Cfg::kABytes, 8192, 1024, 2) : umma_smem_desc(a0 + Cfg::kABytes, 16, 1024, 2);
            constexpr uint32_t kStepA = A_MN ? 2048 : 32, kStepB = B_MN ? 2048 : 32;
            int s = 0;
            uint32_t ph = 0;
            const int kb_tail = num_kb - 1 - ((num_kb - 3) % 3);  // MT = 2 (3 stages, num_kb >= 3): last k-block held in stage 2
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(full_bar + s, ph);
                tc_fence_after();
                const uint64_t da = umma_desc_adv(dA0, s * Cfg::kStageBytes), db = umma_desc_adv(dB0, s * Cfg::kStageBytes);
#pragma unroll
                for (int k = 0; k < GEMM_BK / 16; ++k) {
#pragma unroll
                    for (int t = 0; t < MT; ++t)
                        umma_bf16_ss(tmem_base + t * BN, umma_desc_adv(da, t * (GEMM_BM * GEMM_BK * 2) + k * kStepA),
                                     umma_desc_adv(db, k * kStepB), idesc, (kb | k) ? 1u : 0u);
                }
                umma_commit(empty_bar + s);  // frees the smem stage when these MMAs retire
                if (MT == 2 && kb == kb_tail) umma_commit(tail_bar);
                if (++s == num_stages) {
                    s = 0;
                    ph ^= 1;
                }
            }
            umma_commit(accum_bar);
        }
    } else {
        // ---------------- epilogue warps: TMEM lane quarter = warp % 4, one accumulator row per thread; with 8 warps
        // the two warps of a quarter split the tile's columns in halves
        const int quarter = warp & 3;
        const int half = (warp - 2) >> 2;      // 0 for the 4-warp configurations
        const int c_begin = half * CPW;
        const int row_in_tile = quarter * 32 + lane;
        PhiloxKey key = {0u, 0u, 0u, 0u};
        if (p.drop.state) key = dropout_key(p.drop);

        // phase 0 (overlaps the main loop): per-column parameters of this N tile -> shared memory (coalesced, once),
        const int et = threadIdx.x - 64;  // 0..127 over the four epilogue warps
        for (int i = et; i < BN; i += EW * 32) {
            const bool ok = n0 + i < p.N;
            s_bias[i] = (p.bias && ok) ? __ldg(p.bias + n0 + i) : 0.f;
            if (p.ln_gamma) {
                s_gamma[i] = ok ? __ldg(p.ln_gamma + n0 + i) : 0.f;
                s_beta[i] = ok ? __ldg(p.ln_beta + n0 + i) : 0.f;
            }
        }
        // ... and this thread's residual / mask row(s) -> shared memory, 16 bytes at a time (sub-tile t)
        auto aux_prefetch = [&](int t) {
            const long long r = (long long)m0 + t * GEMM_BM + row_in_tile;
            const bool r_ok = r < p.M;
            const long long mr = p.aux_row_mod > 0 ? (r % p.aux_row_mod) : r;
            const __nv_bfloat16* src = p.aux + (p.aux_mode == 1 ? mr : r) * p.ld_aux + n0;
            const uint32_t dst = smem_u32(aux_stage - t * Cfg::kAuxBytes + row_in_tile * Cfg::kRowBf16);
#pragma unroll 8
            for (int j = c_begin * 4; j < (c_begin + CPW) * 4; ++j) {
                const bool ok = r_ok && (n0 + j * 8 + 8 <= p.N);
                cp_async_16(dst + j * 16, ok ? static_cast<const void*>(src + j * 8) : static_cast<const void*>(p.aux),
                            ok ? 16u : 0u);
            }
            cp_async_commit();
        };
        const bool aux_staged = p.aux_mode && p.vec_ok;
        if (MT == 1 && aux_staged) aux_prefetch(0);

        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");  // parameter slices visible to all epilogue warps
        if (MT == 2 && aux_staged) {  // stage 2 is dead two or three k-blocks before the end: sub-tile 0's rows land meanwhile
            mbar_wait(tail_bar, 0);
            aux_prefetch(0);
        }
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        // The pipeline stages are dead now (every TMA write landed and every MMA retired before accum_bar fired):
        // they become the output staging area (and, for MT = 2, the aux area of sub-tile 1).
        if (MT == 2 && aux_staged) aux_prefetch(1);
        uint8_t* stage_a = smem;
        uint32_t v[32];
        float z[32];
#pragma unroll 1
        for (int t = 0; t < MT; ++t) {
        const long long mt0 = (long long)m0 + t * GEMM_BM;
        const long long row = mt0 + row_in_tile;
        const bool row_ok = row < p.M;
        const long long mod_row = p.aux_row_mod > 0 ? (row % p.aux_row_mod) : row;
        const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + t * BN;
        const uint8_t* aux_row = aux_stage - t * Cfg::kAuxBytes + row_in_tile * Cfg::kRowBf16;
        float* s_stat_t = s_stat + t * (4 * GEMM_BM);
        if (aux_staged) {
            if (MT == 2 && t == 0) cp_async_wait_group<1>();
            else cp_async_wait_all();
        }
        if (MT == 2 && t > 0 && mt0 >= p.M) break;  // (CTA-uniform) the empty second tile of an odd tile count
        if constexpr (!LN) {
            const bool f32out = p.out_f32 != nullptr;
            const int rowb = f32out ? Cfg::kRowF32 : Cfg::kRowBf16;
            uint8_t* my_row = stage_a + row_in_tile * rowb;
#pragma unroll 1
            for (int c = c_begin; c < c_begin + CPW; ++c) {
                const int col0 = n0 + c * 32;
                if (col0 >= p.N) break;  // warp-uniform
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) z[i] = __uint_as_float(v[i]);
                if (row_ok) epi_chunk<BN>(z, p, aux_row, row, mod_row, col0, c, s_bias, key);
                if (f32out) stage_f32(my_row + c * 128, z);
                else stage_bf16(my_row + c * 64, z);
            }
            __syncwarp();
            const uint8_t* stage_warp = stage_a + quarter * 32 * rowb + c_begin * 32 * (f32out ? 4 : 2);
            if (f32out)
                drain_f32<BN, WCOLS>(stage_warp, p.out_f32, p.ld_out, p.M, p.N, mt0 + quarter * 32,
                                     n0 + c_begin * 32, lane, p.vec_ok != 0, p.atomic_out != 0);
            else
                drain_bf16<BN, WCOLS>(stage_warp, p.out_bf16, p.ld_out, p.M, p.N, mt0 + quarter * 32,
                                      n0 + c_begin * 32, lane, p.vec_ok != 0, p.colsum_out);
            if (MT == 2) __syncwarp();  // the staged rows are reused by the next sub-tile
        } else {
            // LayerNorm over the whole row (N == BN == 256, n0 == 0).  The two warps of a lane quarter own 128 columns of
            // the row each.  Two passes over the accumulator IN TMEM (reading it costs ~1 KB/clk per SM, a fraction of
            // what holding 128 values per thread in registers cost in unrolled code: that version was 7 700 SASS
            // instructions executed once per CTA and spent ~45 % of its issue slots waiting for instruction fetch):
            //   pass 1: z = acc + bias + residual -> shifted sums for mean / variance, pre-LN copy staged and drained
            //   pass 2: z again -> (z - mean) * rstd * gamma + beta staged and drained
            // Loops are kept rolled (32 columns per iteration).
            static_assert(BN == 256, "LayerNorm epilogue needs the whole 256-wide row in one tile");
            uint8_t* my_a = stage_a + row_in_tile * Cfg::kRowBf16;
            const long long row0 = mt0 + quarter * 32;
            const int off = c_begin * 64;  // byte offset of this warp's column half inside a staged bf16 row
            float shift = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll 1
            for (int cc = 0; cc < CPW; ++cc) {
                const int c = c_begin + cc;
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
                ln_chunk(z, v, p, aux_row, row, mod_row, c, s_bias, row_ok, key);
                if (cc == 0) shift = z[0];  // sums are taken about a value inside the data: no cancellation in the variance
                float a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float d0 = z[i] - shift, d1 = z[i + 1] - shift;
                    a0 += d0;
                    a1 += d1;
                    q0 = fmaf(d0, d0, q0);
                    q1 = fmaf(d1, d1, q1);
                }
                s1 += a0 + a1;
                s2 += q0 + q1;
                if (p.pre_ln) stage_bf16(my_a + c * 64, z);
            }
            constexpr float kInvHalf = 1.f / (CPW * 32);
            const float m_half = shift + s1 * kInvHalf;           // mean of this warp's 128 columns
            const float m2_half = s2 - s1 * s1 * kInvHalf;        // sum of squared deviations about it
            s_stat_t[half * GEMM_BM + row_in_tile] = m_half;
            s_stat_t[(2 + half) * GEMM_BM + row_in_tile] = m2_half;
            if (p.pre_ln) {
                __syncwarp();
                drain_bf16<BN, WCOLS>(stage_a + quarter * 32 * Cfg::kRowBf16 + off, p.pre_ln, p.ld_out, p.M, p.N, row0,
                                      c_begin * 32, lane, p.vec_ok != 0, nullptr);
            }
            asm volatile("bar.sync 1, %0;" ::"n"(EW * 32) : "memory");
            const float m_other = s_stat_t[(half ^ 1) * GEMM_BM + row_in_tile];
            const float m2_other = s_stat_t[(2 + (half ^ 1)) * GEMM_BM + row_in_tile];
            const float mean = 0.5f * (m_half + m_other);
            const float dm = m_half - m_other;
            // Chan's pairwise combination of two equal halves of n = 128: M2 = M2a + M2b + dm^2 * n / 2
            const float var = (m2_half + m2_other + dm * dm * (0.5f * CPW * 32)) * (1.f / BN);
            const float rstd = rsqrtf(var + p.ln_eps);
            if (row_ok && half == 0) {
                if (p.ln_mean) p.ln_mean[row] = mean;
                if (p.ln_rstd) p.ln_rstd[row] = rstd;
            }
#pragma unroll 1
            for (int cc = 0; cc < CPW; ++cc) {
                const int c = c_begin + cc;
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
                ln_chunk(z, v, p, aux_row, row, mod_row, c, s_bias, row_ok, key);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 g4 = *reinterpret_cast<const float4*>(s_gamma + c * 32 + q * 4);
                    const float4 b4 = *reinterpret_cast<const float4*>(s_beta + c * 32 + q * 4);
                    z[q * 4] = (z[q * 4] - mean) * rstd * g4.x + b4.x;
                    z[q * 4 + 1] = (z[q * 4 + 1] - mean) * rstd * g4.y + b4.y;
                    z[q * 4 + 2] = (z[q * 4 + 2] - mean) * rstd * g4.z + b4.z;
                    z[q * 4 + 3] = (z[q * 4 + 3] - mean) * rstd * g4.w + b4.w;
                }
                stage_bf16(my_a + c * 64, z);  // (this warp's pre-LN rows have been drained: same lanes, program order)
            }
            __syncwarp();
            drain_bf16<BN, WCOLS>(stage_a + quarter * 32 * Cfg::kRowBf16 + off, p.out_bf16, p.ld_out, p.M, p.N, row0,
                                  c_begin * 32, lane, p.vec_ok != 0, nullptr);
            if (MT == 2) __syncwarp();  // the staged rows are reused by the next sub-tile
        }
        }  // sub-tiles
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}

template <int BN, bool A_MN, bool B_MN, bool LN = false, int MT = 1>
static int launch_gemm(const fod_linear_args* a, GemmEpi epi, cudaStream_t st) {
    using Cfg = GemmCfg<BN, MT>;
    CUtensorMap tmA, tmB;
    const long long ldw = a->ldw ? a->ldw : (B_MN ? a->N : a->K);
    if (A_MN) {
        uint64_t dims[2] = {(uint64_t)a->M, (uint64_t)a->K};
        uint64_t str[1] = {(uint64_t)a->ldx * 2};
        uint32_t box[2] = {64, GEMM_BK};
        if (make_tmap_bf16(&tmA, a->x, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm: tmap A (MN-major)");
    } else {
        uint64_t dims[2] = {(uint64_t)a->K, (uint64_t)a->M};
        uint64_t str[1] = {(uint64_t)a->ldx * 2};
        uint32_t box[2] = {GEMM_BK, GEMM_BM};
        if (make_tmap_bf16(&tmA, a->x, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm: tmap A");
    }
    if (B_MN) {
        uint64_t dims[2] = {(uint64_t)a->N, (uint64_t)a->K};
        uint64_t str[1] = {(uint64_t)ldw * 2};
        uint32_t box[2] = {64, GEMM_BK};
        if (make_tmap_bf16(&tmB, a->w, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm: tmap B (MN-major)");
    } else {
        uint64_t dims[2] = {(uint64_t)a->K, (uint64_t)a->N};
        uint64_t str[1] = {(uint64_t)ldw * 2};
        uint32_t box[2] = {GEMM_BK, (uint32_t)BN};
        if (make_tmap_bf16(&tmB, a->w, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm: tmap B");
    }
    const bool aux = epi.aux_mode != 0;
    // stage count: the pipeline stages double as output staging, so they must hold the staged tile
    int stages = BN == 256 ? (aux ? 3 : 4) : (BN == 128 ? 4 : 3);
    const int total_kb = (a->K + GEMM_BK - 1) / GEMM_BK;
    const int mt = (a->M + GEMM_BM * MT - 1) / (GEMM_BM * MT), nt = (a->N + BN - 1) / BN;
    int splits = 1;
    if (a->split_k > 1) splits = a->split_k;
    else if (a->split_k < 0) {  // auto: fill the 148 SMs, at least two k-blocks per split
        splits = (148 + mt * nt - 1) / (mt * nt);
        if (BN == 256 && mt * nt <= 148) splits = 148 / (mt * nt);  // one CTA per SM at this tile width: no partial second wave
        if (splits > total_kb / 2) splits = total_kb / 2;
        if (splits < 1) splits = 1;
    }
    if (splits > total_kb) splits = total_kb;
    epi.kb_per_split = (total_kb + splits - 1) / splits;
    splits = (total_kb + epi.kb_per_split - 1) / epi.kb_per_split;  // no empty split
    // Short contractions (K <= 256: every projection of the 256-wide model) need no deep pipeline: with 2 stages two
    // CTAs fit one SM (smem and the 512 TMEM columns), so one CTA's epilogue overlaps the other's main loop.  The
    // stages must still hold the staged output tile (and the pre-LN copy for the LayerNorm epilogue).
    {
        const int need = a->out_f32 ? GEMM_BM * Cfg::kRowF32 : GEMM_BM * Cfg::kRowBf16;  // (pre-LN copy and output share it)
        int min_stages = (need + Cfg::kStageBytes - 1) / Cfg::kStageBytes;
        if (min_stages < 2) min_stages = 2;
        const int want = epi.kb_per_split < min_stages ? min_stages : epi.kb_per_split;
        if (want < stages) stages = want;
    }
    if (MT == 2) {  // (gemm_two_tiles() admits only what this layout holds: bf16 output, no split, >= 3 k-blocks)
        if (splits != 1 || a->out_f32 || total_kb < Cfg::kMt2Stages) return fod_fail(FOD_E_ARG, "gemm: two-tile CTAs misrouted");
        stages = Cfg::kMt2Stages;
    }
    bool memset_before = false;
    if (a->split_k < 0 && !a->atomic_out) {  // overwrite semantics: plain stores, or zero-fill + reductions
        if (splits > 1) {
            memset_before = true;
            if (a->ld_out != a->N) return fod_fail(FOD_E_ARG, "gemm: auto split-K needs a dense fp32 output");
            cudaMemsetAsync(a->out_f32, 0, sizeof(float) * (size_t)a->M * a->N, st);
            epi.atomic_out = 1;
        }
    }
    epi.num_stages = stages;
    const int smem_bytes = Cfg::smem_bytes(stages, aux);
    static bool attr_set = false;
    if (!attr_set) {
        const int a1 = Cfg::smem_bytes(BN == 256 ? 3 : (BN == 128 ? 4 : 3), true), a2 = Cfg::smem_bytes(BN == 64 ? 3 : 4, false);
        cudaFuncSetAttribute(gemm_bf16_kernel<BN, A_MN, B_MN, LN, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             a1 > a2 ? a1 : a2);
        attr_set = true;
    }
    dim3 grid(nt, mt, splits);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(GemmThreads<BN>::kThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (fod_pdl_enabled() && !memset_before) ? 1 : 0;  // a memset node cannot be a programmatic predecessor
    cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<BN, A_MN, B_MN, LN, MT>, tmA, tmB, epi);
    return fod_check_launch("gemm_bf16_kernel");
}

// =====================================================================================================================
// Persistent weight-stationary kernel for the short contractions (K <= 256: every projection OUT of the 256-wide model —
// FFN up-projection and its dX twin dH = dU·W2, the fused QKV projection).  The one-tile-per-CTA kernel above spends a
// K = 256 tile mostly on things that are not the 16 MMAs: CTA launch, barrier / TMEM set-up, the first TMA round trip,
// 128 KB of W re-fetched from L2 for every tile, and an epilogue nothing overlaps (one CTA per SM at BN = 256).  Here
//   * one CTA per SM lives for the whole GEMM and keeps ITS 256-column slab of W (<= 128 KB) in shared memory: only
//     the 64 KB activation tiles stream (3-stage ring of 16 KB k-blocks) — a third of the L2 -> SM bytes per tile;
//   * the accumulator is double-buffered in TMEM (2 x 256 columns): the MMAs of tile i+1 run while the eight epilogue
//     warps drain tile i, so bias / ReLU / mask / dropout arithmetic and the stores are off the critical path;
//   * the epilogue works on 32-column chunks through small per-warp staging buffers (coalesced 64-byte row segments
//     out, coalesced cp.async of the residual / mask chunk in, one chunk ahead), and bias-gradient column sums stay
//     in registers for the CTA's whole column slab: one atomic per column per warp instead of one per tile.
// CTA c keeps W slab c / ctas_per_n and takes the M tiles first, first + ctas_per_n, ... of that slab.
constexpr int WS_BN = 256;
constexpr int WS_EPI_WARPS = 8;
constexpr int WS_THREADS = 64 + 32 * WS_EPI_WARPS;
constexpr int WS_A_STAGES = 3;
constexpr int WS_A_BYTES = GEMM_BM * GEMM_BK * 2;           // 16 KB: one k-block of the activation tile
constexpr int WS_B_KB_BYTES = WS_BN * GEMM_BK * 2;          // 32 KB: one k-block of the resident weight slab
constexpr int WS_B_BYTES = 4 * WS_B_KB_BYTES;               // K <= 256
constexpr int WS_ROWB = 80;                                 // staged row: 32 bf16 + 16 B padding (conflict-free both ways)
constexpr int WS_XBUF = 32 * WS_ROWB;
constexpr int WS_X_BYTES = WS_EPI_WARPS * 2 * WS_XBUF;      // two chunk buffers per epilogue warp
constexpr int WS_CS_BYTES = WS_EPI_WARPS * 128 * 4;            // per-warp column-sum accumulators (bias gradients)
constexpr int WS_SMEM = 1024 + WS_B_BYTES + WS_A_STAGES * WS_A_BYTES + WS_X_BYTES + 256 + WS_BN * 4 + WS_CS_BYTES;


__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 u;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "r"(addr) : "memory");
    return u;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& u) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}

// residual / mask chunk (32 rows x 32 columns of this warp) -> staging buffer, 4 lanes per 64-byte row segment
__device__ __forceinline__ void ws_aux_load(uint32_t xb, const GemmEpi& p, long long row0, int col0, int lane) {
    const int piece = lane & 3, r0 = lane >> 2;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + 8 * i;
        const long long grow = row0 + r;
        const bool ok = grow < p.M;
        const long long srow = (p.aux_mode == 1 && p.aux_row_mod > 0) ? (grow % p.aux_row_mod) : grow;
        const __nv_bfloat16* src = p.aux + srow * p.ld_aux + col0 + piece * 8;
        cp_async_16(xb + r * WS_ROWB + piece * 16, ok ? static_cast<const void*>(src) : static_cast<const void*>(p.aux),
                    ok ? 16u : 0u);
    }
}

template <bool DROP>
__device__ __forceinline__ void ws_epi_chunk(float (&z)[32], const GemmEpi& p, uint32_t xrow, long long row, int col0,
                                             const float* bias32, const PhiloxKey& key) {
    if (p.bias) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float4 b4 = *reinterpret_cast<const float4*>(bias32 + q * 4);
            z[q * 4] += b4.x;
            z[q * 4 + 1] += b4.y;
            z[q * 4 + 2] += b4.z;
            z[q * 4 + 3] += b4.w;
        }
    }
    if (p.aux_mode == 1) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float f[8];
            unpack8(lds128(xrow + q * 16), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) z[q * 8 + j] += f[j];
        }
    }
    if (p.relu) {
#pragma unroll
        for (int i = 0; i < 32; ++i) z[i] = fmaxf(z[i], 0.f);
    }
    if (p.aux_mode == 2) {
        const float ms = p.mask_scale;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float f[8];
            unpack8(lds128(xrow + q * 16), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) z[q * 8 + j] = (f[j] > 0.f) ? z[q * 8 + j] * ms : 0.f;
        }
    }
    if constexpr (DROP) drop_chunk(z, key, p.drop, row, p.N, col0);
}

__device__ __forceinline__ void ws_stage(uint32_t xrow, const float (&z)[32]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(z[q * 8 + 0], z[q * 8 + 1]);
        u.y = pack_bf16x2(z[q * 8 + 2], z[q * 8 + 3]);
        u.z = pack_bf16x2(z[q * 8 + 4], z[q * 8 + 5]);
        u.w = pack_bf16x2(z[q * 8 + 6], z[q * 8 + 7]);
        sts128(xrow + q * 16, u);
    }
}

// staged chunk -> global memory: one warp instruction stores 8 row segments of 64 bytes.  The four shared-memory reads
// are issued together, the stores are predicated (no branch per row).
__device__ __forceinline__ void ws_drain(uint32_t xb, const GemmEpi& p, long long row0, int col0, int lane, float* cs32) {
    const int piece = lane & 3, r0 = lane >> 2;
    uint4 u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) u[i] = lds128(xb + (r0 + 8 * i) * WS_ROWB + piece * 16);
    __nv_bfloat16* dst = p.out_bf16 + (row0 + r0) * p.ld_out + col0 + piece * 8;
    const long long left = p.M - (row0 + r0);  // rows r0 + 8 i are valid while 8 i < left
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (8 * i < left) {
            *reinterpret_cast<uint4*>(dst + (long long)(8 * i) * p.ld_out) = u[i];
        } else {
            u[i] = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    if (p.colsum_out) {  // column sums of the stored bf16 values: 4 rows per lane, then the 8 lanes that share a piece
        float cs[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) cs[j] = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float f[8];
            unpack8(u[i], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) cs[j] += f[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = cs[j];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (lane < 4) cs32[lane * 8 + j] += v;  // lane-private slots of this warp's accumulator row
        }
    }
}

template <bool B_MN, bool DROP>
__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_ws_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmEpi p, int mt,
               int ctas_per_n) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sB = smem;
    uint8_t* sA = sB + WS_B_BYTES;
    uint8_t* sX = sA + WS_A_STAGES * WS_A_BYTES;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(sX + WS_X_BYTES);
    uint64_t* a_empty = a_full + WS_A_STAGES;
    uint64_t* b_full = a_empty + WS_A_STAGES;
    uint64_t* t_full = b_full + 1;
    uint64_t* t_empty = t_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
    float* s_bias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(a_full) + 256);
    float* s_cs = s_bias + WS_BN;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = (int)(blockIdx.x / ctas_per_n) * WS_BN;
    const int first = blockIdx.x % ctas_per_n;
    const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < WS_A_STAGES; ++s) {
            mbar_init(a_full + s, 1);
            mbar_init(a_empty + s, 1);
        }
        mbar_init(b_full, 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(t_full + b, 1);
            mbar_init(t_empty + b, WS_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(b_full, (uint32_t)num_kb * WS_B_KB_BYTES);
            for (int kb = 0; kb < num_kb; ++kb) {
                uint8_t* sb = sB + kb * WS_B_KB_BYTES;
                if constexpr (B_MN) {
#pragma unroll
                    for (int c = 0; c < WS_BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, b_full, n0 + c * 64, kb * GEMM_BK);
                } else {
                    tma_load_2d(sb, &tmB, b_full, kb * GEMM_BK, n0);
                }
            }
            int s = 0;
            uint32_t ph = 0;
            for (int t = first; t < mt; t += ctas_per_n) {
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(a_empty + s, ph ^ 1);
                    mbar_arrive_expect_tx(a_full + s, WS_A_BYTES);
                    tma_load_2d(sA + s * WS_A_BYTES, &tmA, a_full + s, kb * GEMM_BK, t * GEMM_BM);
                    if (++s == WS_A_STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, WS_BN, 0, B_MN ? 1 : 0);
            const uint64_t dA0 = umma_smem_desc(smem_u32(sA), 16, 1024, 2);
            const uint64_t dB0 = B_MN ? umma_smem_desc(smem_u32(sB), 8192, 1024, 2) : umma_smem_desc(smem_u32(sB), 16, 1024, 2);
            constexpr uint32_t kStepB = B_MN ? 2048 : 32;
            mbar_wait(b_full, 0);
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int t = first; t < mt; t += ctas_per_n, ++it) {
                const int buf = it & 1;
                mbar_wait(t_empty + buf, ((it >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator buffer
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * WS_BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(a_full + s, ph);
                    tc_fence_after();
                    const uint64_t da = umma_desc_adv(dA0, s * WS_A_BYTES), db = umma_desc_adv(dB0, kb * WS_B_KB_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; ++k)
                        umma_bf16_ss(d_tmem, umma_desc_adv(da, k * 32), umma_desc_adv(db, k * kStepB), idesc, (kb | k) ? 1u : 0u);
                    umma_commit(a_empty + s);
                    if (++s == WS_A_STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
                umma_commit(t_full + buf);
            }
        }
    } else {
        const int quarter = warp & 3;        // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;    // which 128 columns of the tile
        const uint32_t X = smem_u32(sX + (warp - 2) * 2 * WS_XBUF);
        for (int i = threadIdx.x - 64; i < WS_BN; i += WS_EPI_WARPS * 32) s_bias[i] = p.bias ? __ldg(p.bias + n0 + i) : 0.f;
        asm volatile("bar.sync 1, %0;" ::"n"(WS_EPI_WARPS * 32) : "memory");
        PhiloxKey key = {0u, 0u, 0u, 0u};
        if constexpr (DROP) key = dropout_key(p.drop);
        float* cs_w = s_cs + (warp - 2) * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i) cs_w[lane * 4 + i] = 0.f;
        __syncwarp();
        const int colw = n0 + half * 128;  // first column of this warp
        int it = 0;
        for (int t = first; t < mt; t += ctas_per_n, ++it) {
            const int buf = it & 1;
            const long long row0 = (long long)t * GEMM_BM + quarter * 32;
            const long long row = row0 + lane;
            const bool row_ok = row < p.M;
            if (p.aux_mode) ws_aux_load(X, p, row0, colw, lane);
            cp_async_commit();
            mbar_wait(t_full + buf, (it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * WS_BN + half * 128 + (uint32_t(quarter * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                const uint32_t xb = X + (c & 1) * WS_XBUF;
                if (c < 3 && p.aux_mode) ws_aux_load(X + ((c + 1) & 1) * WS_XBUF, p, row0, colw + (c + 1) * 32, lane);
                cp_async_commit();
                uint32_t v[32];
                float z[32];
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
                if (c == 3) {  // the whole accumulator is in registers: hand the buffer back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(t_empty + buf);
                }
                cp_async_wait_group<1>();
                __syncwarp();
#pragma unroll
                for (int i = 0; i < 32; ++i) z[i] = __uint_as_float(v[i]);
                const uint32_t xrow = xb + lane * WS_ROWB;
                if (row_ok) ws_epi_chunk<DROP>(z, p, xrow, row, colw + c * 32, s_bias + half * 128 + c * 32, key);
                ws_stage(xrow, z);
                __syncwarp();
                ws_drain(xb, p, row0, colw + c * 32, lane, cs_w + c * 32);
                __syncwarp();
            }
        }
        if (p.colsum_out) {
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i) atomicAdd(p.colsum_out + colw + i * 32 + lane, cs_w[i * 32 + lane]);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// shapes the persistent kernel takes: see the dispatch in fod_linear_fwd
static bool gemm_ws_eligible(const fod_linear_args* a, const GemmEpi& e) {
    static const int enabled = [] { const char* v = getenv("FOD_GEMM_WS"); return v ? atoi(v) : 1; }();
    if (!enabled) return false;
    if (a->a_mn || a->ln_gamma || !a->out_bf16 || a->col_scale || a->split_k > 1 || a->split_k < 0 || !e.vec_ok) return false;
    if (a->K > 256 || a->K % GEMM_BK || a->N % WS_BN) return false;
    const long long tiles = (long long)((a->M + GEMM_BM - 1) / GEMM_BM) * (a->N / WS_BN);
    return tiles >= 2 * 148;  // at least two tiles per SM: something to overlap
}

template <bool B_MN, bool DROP>
static int launch_gemm_ws(const fod_linear_args* a, const GemmEpi& epi, cudaStream_t st) {
    CUtensorMap tmA, tmB;
    const long long ldw = a->ldw ? a->ldw : (B_MN ? a->N : a->K);
    {
        uint64_t dims[2] = {(uint64_t)a->K, (uint64_t)a->M};
        uint64_t str[1] = {(uint64_t)a->ldx * 2};
        uint32_t box[2] = {GEMM_BK, GEMM_BM};
        if (make_tmap_bf16(&tmA, a->x, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm(ws): tmap A");
    }
    if (B_MN) {
        uint64_t dims[2] = {(uint64_t)a->N, (uint64_t)a->K};
        uint64_t str[1] = {(uint64_t)ldw * 2};
        uint32_t box[2] = {64, GEMM_BK};
        if (make_tmap_bf16(&tmB, a->w, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm(ws): tmap B (MN-major)");
    } else {
        uint64_t dims[2] = {(uint64_t)a->K, (uint64_t)a->N};
        uint64_t str[1] = {(uint64_t)ldw * 2};
        uint32_t box[2] = {GEMM_BK, (uint32_t)WS_BN};
        if (make_tmap_bf16(&tmB, a->w, 2, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "gemm(ws): tmap B");
    }
    const int mt = (a->M + GEMM_BM - 1) / GEMM_BM, nt = a->N / WS_BN;
    int cpn = 148 / nt;
    if (cpn < 1) cpn = 1;
    if (cpn > mt) cpn = mt;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(gemm_ws_kernel<B_MN, DROP>, cudaFuncAttributeMaxDynamicSharedMemorySize, WS_SMEM);
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nt * cpn);
    cfg.blockDim = dim3(WS_THREADS);
    cfg.dynamicSmemBytes = WS_SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = fod_pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, gemm_ws_kernel<B_MN, DROP>, tmA, tmB, epi, mt, cpn);
    return fod_check_launch("gemm_ws_kernel");
}

// Two 128-row tiles per CTA (GemmCfg<256, 2>) when the 128-row tiling leaves a partial last wave that the 256-row
// tiling removes.  A 256-row CTA is charged as 1.9 one-tile CTAs (28.8 us for one wave of them against 30.1 / 2 us per
// one-tile wave, plain epilogue): 22 400 rows, 175 tiles = 2 waves -> 88 CTAs = 1 wave qualifies; 44 800 rows
// (350 tiles = 3 waves -> 175 CTAs = 2 waves) does not, and measured no different in the stage-2 step.
// Measured at 22 400 rows (us, one tile -> two tiles per CTA): LayerNorm epilogue K = 2048 41.3 -> 38.1, with dropout and
// saved statistics 53.3 -> 48.4, dX + residual 37.3 -> 33.9, plain 30.1 -> 28.8; K = 256: 11.0 -> 9.8, LayerNorm 18.6 -> 17.9.
// FOD_GEMM_MT2=0 disables the path, =1 keeps it to contractions of >= 8 k-blocks.
static bool gemm_two_tiles(const fod_linear_args* a, const GemmEpi& e) {
    static const int mode = [] { const char* v = getenv("FOD_GEMM_MT2"); return v ? atoi(v) : 2; }();
    if (mode == 0 || a->a_mn || a->N > 256 || a->N <= 128 || a->out_f32 || a->split_k || a->atomic_out || !e.vec_ok) return false;
    const int total_kb = (a->K + GEMM_BK - 1) / GEMM_BK;
    if (total_kb < (mode == 2 ? 3 : 8)) return false;
    const int mt1 = (a->M + GEMM_BM - 1) / GEMM_BM, mt2 = (mt1 + 1) / 2;
    return 1.9 * ((mt2 + 147) / 148) < 1.0 * ((mt1 + 147) / 148);
}

template <int BN>
static int launch_gemm_major(const fod_linear_args* a, const GemmEpi& e, cudaStream_t st) {
    if constexpr (BN == 256) {
        const bool two = gemm_two_tiles(a, e);
        if (a->ln_gamma) {
            if (a->a_mn || a->b_mn) return fod_fail(FOD_E_ARG, "gemm: the LayerNorm epilogue takes K-major operands");
            if (two) return launch_gemm<256, false, false, true, 2>(a, e, st);
            return launch_gemm<256, false, false, true>(a, e, st);
        }
        if (two) return a->b_mn ? launch_gemm<256, false, true, false, 2>(a, e, st) : launch_gemm<256, false, false, false, 2>(a, e, st);
    }
    if (a->a_mn && a->b_mn) return launch_gemm<BN, true, true>(a, e, st);
    if (a->b_mn) return launch_gemm<BN, false, true>(a, e, st);
    if (a->a_mn) return fod_fail(FOD_E_ARG, "gemm: A MN-major with B K-major is not instantiated");
    return launch_gemm<BN, false, false>(a, e, st);
}

// Largest N tile that still yields >= half a wave of CTAs; small problems take the narrowest tile (more CTAs, a
// 4x shorter epilogue per CTA).  LayerNorm needs the whole 256-wide row in one CTA.
static int pick_bn(const fod_linear_args* a) {
    if (a->ln_gamma) return 256;
    if (a->N <= 64) return 64;
    static const int forced = [] { const char* e = getenv("FOD_GEMM_BN"); return e ? atoi(e) : 0; }();  // tuning probe
    if (forced == 64 || forced == 128 || (forced == 256 && a->N > 128)) return forced;
    const int mt = (a->M + GEMM_BM - 1) / GEMM_BM;
    // Contractions the library may split (weight gradients: K = token rows): the SMs are filled by split-K, so take the
    // widest tile as long as every split keeps >= 6 k-blocks of work — an M128 x K16 MMA costs ~52 clk for any N <= 64
    // and 132 clk for N = 256 (csrc/microbench.cu), the A tile is fetched once per 256 instead of per 64 columns, and
    // 148 / tiles splits make exactly one wave.  Measured (us, before -> after): K = 22 400: dW 2048x256 45.7 -> 22.9,
    // 256x2048 48.6 -> 25.1, 768x256 19.8 -> 11.8; K = 4 096: 2048x256 11.2 -> 9.1; below the threshold the narrow tile wins
    // (K = 4 096: 768x256 7.1 vs 7.7, 256x256 4.7 vs 6.7).
    if ((a->split_k < 0 || a->atomic_out) && a->split_k <= 1 && a->N >= 256) {
        const long long total_kb = (a->K + GEMM_BK - 1) / GEMM_BK, tiles = (long long)mt * ((a->N + 255) / 256);
        if (tiles <= 148 && total_kb * tiles >= 6 * 148) return 256;
    }
    const int cands[3] = {256, 128, 64};
    for (int i = 0; i < 3; ++i) {
        const int bn = cands[i];
        if (bn == 256 && a->N <= 128) continue;
        if (mt * ((a->N + bn - 1) / bn) >= 74) return bn;
    }
    return 64;
}

}  // namespace fod

extern "C" int fod_linear_fwd(const fod_linear_args* a, void* stream) {
    using namespace fod;
    if (!a || !a->x || !a->w || a->M <= 0 || a->N <= 0 || a->K <= 0) return fod_fail(FOD_E_ARG, "linear: null/empty");
    if (a->ldx % 8 || a->ldw % 8) return fod_fail(FOD_E_ARG, "linear: ldx/ldw must be multiples of 8");
    if (!a->a_mn && (a->K % 8)) return fod_fail(FOD_E_ARG, "linear: K must be a multiple of 8");
    if (a->a_mn && (a->M % 8)) return fod_fail(FOD_E_ARG, "linear: M must be a multiple of 8 for MN-major A");
    if (a->b_mn ? (a->N % 8) : (a->K % 8)) return fod_fail(FOD_E_ARG, "linear: W inner extent must be a multiple of 8");
    if ((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->w)) & 15)
        return fod_fail(FOD_E_ARG, "linear: x/w must be 16-byte aligned");
    if (a->ln_gamma && (a->N != 256 || !a->ln_beta || a->relu || a->split_k > 1 || a->relu_mask || a->out_f32 || !a->out_bf16))
        return fod_fail(FOD_E_ARG, "linear: LayerNorm epilogue needs N == 256, beta, bf16 output, no relu/mask/split");
    if (!a->out_bf16 && !a->out_f32) return fod_fail(FOD_E_ARG, "linear: no output");
    if (a->out_bf16 && a->out_f32) return fod_fail(FOD_E_ARG, "linear: one output (bf16 or fp32)");
    if (a->residual && a->relu_mask) return fod_fail(FOD_E_ARG, "linear: residual and relu_mask are exclusive");
    if ((a->split_k > 1 || a->split_k < 0) && (!a->out_f32 || a->relu || a->bias || a->residual || a->relu_mask))
        return fod_fail(FOD_E_ARG, "linear: split_k needs fp32 output and a plain epilogue");
    if (a->split_k > 1 && !a->atomic_out) return fod_fail(FOD_E_ARG, "linear: explicit split_k needs atomic_out");
    if (a->atomic_out && !a->out_f32) return fod_fail(FOD_E_ARG, "linear: atomic_out needs out_f32");
    if (a->colsum_out && (!a->out_bf16 || a->ln_gamma)) return fod_fail(FOD_E_ARG, "linear: colsum_out needs a plain bf16 output");
    GemmEpi e;
    e.bias = a->bias;
    e.col_scale = a->col_scale;
    e.atomic_out = a->atomic_out;
    e.kb_per_split = 0;
    e.num_stages = 0;
    e.ln_gamma = a->ln_gamma;
    e.ln_beta = a->ln_beta;
    e.out_bf16 = reinterpret_cast<__nv_bfloat16*>(a->out_bf16);
    e.out_f32 = a->out_f32;
    e.pre_ln = reinterpret_cast<__nv_bfloat16*>(a->pre_ln_bf16);
    e.ln_mean = a->ln_mean;
    e.ln_rstd = a->ln_rstd;
    e.colsum_out = a->colsum_out;
    e.ld_out = a->ld_out;
    e.M = a->M;
    e.N = a->N;
    e.K = a->K;
    e.relu = a->relu;
    e.ln_eps = a->ln_eps;
    e.mask_scale = a->mask_scale != 0.f ? a->mask_scale : 1.f;
    e.drop.state = nullptr;
    e.drop.call = 0;
    e.drop.thresh = 0;
    e.drop.scale = 1.f;
    if (a->drop.rng_state && a->drop.p > 0.f) {
        if (!(a->drop.p < 1.f) || a->N % 8) return fod_fail(FOD_E_ARG, "linear: dropout needs 0 <= p < 1 and N % 8 == 0");
        if (a->out_f32 || a->relu_mask || (a->residual && !a->ln_gamma))
            return fod_fail(FOD_E_ARG, "linear: dropout goes with a bf16 output, and with a residual only under the LayerNorm epilogue");
        e.drop.state = reinterpret_cast<const unsigned long long*>(a->drop.rng_state);
        e.drop.call = (unsigned int)a->drop.call;
        e.drop.thresh = (unsigned int)(a->drop.p * 65536.f + 0.5f);
        e.drop.scale = 1.f / (1.f - a->drop.p);
    }
    e.aux = nullptr;
    e.ld_aux = 0;
    e.aux_mode = 0;
    e.aux_row_mod = a->res_row_mod;  // also indexes col_scale
    if (a->residual) {
        e.aux = reinterpret_cast<const __nv_bfloat16*>(a->residual);
        e.ld_aux = a->ld_res;
        e.aux_mode = 1;
    } else if (a->relu_mask) {
        e.aux = reinterpret_cast<const __nv_bfloat16*>(a->relu_mask);
        e.ld_aux = a->ld_out;
        e.aux_mode = 2;
    }
    // 16-byte vector paths: every row start and every 8-column (bf16) / 4-column (fp32) chunk must be aligned
    bool vec = (a->N % 8 == 0);
    if (a->out_bf16) vec = vec && (a->ld_out % 8 == 0) && !(reinterpret_cast<uintptr_t>(a->out_bf16) & 15);
    if (a->out_f32) vec = vec && (a->ld_out % 4 == 0) && !(reinterpret_cast<uintptr_t>(a->out_f32) & 15);
    if (a->pre_ln_bf16) vec = vec && !(reinterpret_cast<uintptr_t>(a->pre_ln_bf16) & 15);
    if (e.aux) vec = vec && (e.ld_aux % 8 == 0) && !(reinterpret_cast<uintptr_t>(e.aux) & 15);
    e.vec_ok = vec ? 1 : 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (gemm_ws_eligible(a, e)) {
        if (e.drop.state) return a->b_mn ? launch_gemm_ws<true, true>(a, e, st) : launch_gemm_ws<false, true>(a, e, st);
        return a->b_mn ? launch_gemm_ws<true, false>(a, e, st) : launch_gemm_ws<false, false>(a, e, st);
    }
    switch (pick_bn(a)) {
        case 256: return launch_gemm_major<256>(a, e, st);
        case 128: return launch_gemm_major<128>(a, e, st);
        default: return launch_gemm_major<64>(a, e, st);
    }
}
