// Y = epilogue(X · Wᵀ): warp-specialised TMA -> tcgen05.mma -> TMEM -> fused epilogue GEMM for sm_100a.
//   warp 0: TMA producer (A = X tile 128x64, B = W tile BNx64, SWIZZLE_128B, K-major both)
//   warp 1: TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=BN, K=16 per instruction)
//   warps 2..5: epilogue, one accumulator row per thread (tcgen05.ld 32x32b), bias / residual / ReLU /
//               LayerNorm over the full 256-wide row without leaving the CTA.
// Roofline: tensor-bound for the encoder FFN (K=256/2048, M>=5600); weight/launch bound for the decoder.
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"

namespace fod {

struct GemmEpi {
    const float* bias;
    const __nv_bfloat16* residual;
    const __nv_bfloat16* relu_mask;  // (M,N), ld_out: output is zeroed where mask <= 0 (ReLU backward)
    const float* col_scale;          // optional (res_row_mod or M, N) fp32 multiplier applied after bias: y = (acc+b)*s
    int atomic_out;                  // 1: red.add into out_f32 (split-K / gradient accumulation)
    int kb_per_split;
    const float* ln_gamma;
    const float* ln_beta;
    __nv_bfloat16* out_bf16;
    float* out_f32;
    __nv_bfloat16* pre_ln;
    float* ln_mean;
    float* ln_rstd;
    long long ld_res, ld_out;
    int M, N, K;
    int res_row_mod;
    int relu;
    float ln_eps;
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 192;

template <int BN>
struct GemmCfg {
    static constexpr int kABytes = GEMM_BM * GEMM_BK * 2;
    static constexpr int kBBytes = BN * GEMM_BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
    static constexpr int kTmemCols = BN < 32 ? 32 : BN;
};

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]),
        "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]),
        "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// add bias and residual to 32 accumulator columns [col0, col0+32) of one row
__device__ __forceinline__ void epi_bias_res(float (&z)[32], const GemmEpi& p, long long res_row, int col0) {
    if (p.bias) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (col0 + i < p.N) z[i] += __ldg(p.bias + col0 + i);
    }
    if (p.col_scale) {
        const float* sc = p.col_scale + res_row * (long long)p.N + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (col0 + i < p.N) z[i] *= __ldg(sc + i);
    }
    if (p.residual) {
        const __nv_bfloat16* r = p.residual + res_row * p.ld_res + col0;
        if (col0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(r) & 15) == 0)) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint4 u = __ldg(reinterpret_cast<const uint4*>(r) + q);
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float2 f = __bfloat1622float2(h[j]);
                    z[q * 8 + 2 * j] += f.x;
                    z[q * 8 + 2 * j + 1] += f.y;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) z[i] += __bfloat162float(r[i]);
        }
    }
}

__device__ __forceinline__ void epi_store_bf16(__nv_bfloat16* dst, const float (&z)[32], int ncols_valid) {
    if (ncols_valid >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint4 u;
            u.x = pack_bf16x2(z[q * 8 + 0], z[q * 8 + 1]);
            u.y = pack_bf16x2(z[q * 8 + 2], z[q * 8 + 3]);
            u.z = pack_bf16x2(z[q * 8 + 4], z[q * 8 + 5]);
            u.w = pack_bf16x2(z[q * 8 + 6], z[q * 8 + 7]);
            reinterpret_cast<uint4*>(dst)[q] = u;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (i < ncols_valid) dst[i] = __float2bfloat16(z[i]);
    }
}
__device__ __forceinline__ void epi_store_f32(float* dst, const float (&z)[32], int ncols_valid) {
    if (ncols_valid >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
        for (int q = 0; q < 8; ++q)
            reinterpret_cast<float4*>(dst)[q] = make_float4(z[q * 4], z[q * 4 + 1], z[q * 4 + 2], z[q * 4 + 3]);
    } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
            if (i < ncols_valid) dst[i] = z[i];
    }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmEpi p) {
    using Cfg = GemmCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* accum_bar = empty_bar + Cfg::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN;
    const int m0 = blockIdx.y * GEMM_BM;
    const int total_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
    const int kb_begin = blockIdx.z * p.kb_per_split;
    const int kb_end = min(total_kb, kb_begin + p.kb_per_split);
    const int num_kb = kb_end - kb_begin;  // >= 1 by construction of the grid

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(full_bar + s, 1);
            mbar_init(empty_bar + s, 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % Cfg::kStages;
                const uint32_t ph = (kb / Cfg::kStages) & 1;
                mbar_wait(empty_bar + s, ph ^ 1);
                uint8_t* sa = smem + s * Cfg::kStageBytes;
                uint8_t* sb = sa + Cfg::kABytes;
                mbar_arrive_expect_tx(full_bar + s, Cfg::kStageBytes);
                const int k0 = (kb_begin + kb) * GEMM_BK;
                if constexpr (A_MN) {  // memory is (K, M): 64x64 boxes, one per 64-wide M chunk
#pragma unroll
                    for (int c = 0; c < GEMM_BM / 64; ++c) tma_load_2d(sa + c * 8192, &tmA, full_bar + s, m0 + c * 64, k0);
                } else {
                    tma_load_2d(sa, &tmA, full_bar + s, k0, m0);
                }
                if constexpr (B_MN) {
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, full_bar + s, n0 + c * 64, k0);
                } else {
                    tma_load_2d(sb, &tmB, full_bar + s, k0, n0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % Cfg::kStages;
                const uint32_t ph = (kb / Cfg::kStages) & 1;
                mbar_wait(full_bar + s, ph);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * Cfg::kStageBytes);
                const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
                for (int k = 0; k < GEMM_BK / 16; ++k) {
                    // K-major: 128-byte rows of 64 k, 8-row groups 1024 B apart.  MN-major: [64 k][64 mn] slabs of
                    // 8 KB (LBO = next 64 mn), 8-k groups 1024 B apart (SBO), 16 k per instruction = 2048 B.
                    const uint64_t da = A_MN ? umma_smem_desc(a_addr + k * 2048, 8192, 1024, 2)
                                             : umma_smem_desc(a_addr + k * 32, 16, 1024, 2);
                    const uint64_t db = B_MN ? umma_smem_desc(b_addr + k * 2048, 8192, 1024, 2)
                                             : umma_smem_desc(b_addr + k * 32, 16, 1024, 2);
                    umma_bf16_ss(tmem_base, da, db, idesc, (kb | k) ? 1u : 0u);
                }
                umma_commit(empty_bar + s);  // frees the smem stage when these MMAs retire
            }
            umma_commit(accum_bar);
        }
    } else {
        // ---------------- epilogue: warps 2..5, lane quarter = warp % 4
        const int quarter = warp & 3;
        const int row_in_tile = quarter * 32 + lane;
        const long long row = (long long)m0 + row_in_tile;
        const bool row_ok = row < p.M;
        const long long res_row = p.res_row_mod > 0 ? (row % p.res_row_mod) : row;
        const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16);
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        uint32_t v[32];
        float z[32];
        if (p.ln_gamma == nullptr) {
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                const int col0 = n0 + c * 32;
                if (col0 >= p.N) break;  // warp-uniform
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) z[i] = __uint_as_float(v[i]);
                if (row_ok) {
                    epi_bias_res(z, p, res_row, col0);
                    if (p.relu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) z[i] = fmaxf(z[i], 0.f);
                    }
                    const int nv = p.N - col0;
                    if (p.relu_mask) {
                        const __nv_bfloat16* mk = p.relu_mask + row * p.ld_out + col0;
                        if (nv >= 32 && ((reinterpret_cast<uintptr_t>(mk) & 15) == 0)) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint4 u = __ldg(reinterpret_cast<const uint4*>(mk) + q);
                                const __nv_bfloat16* hm = reinterpret_cast<const __nv_bfloat16*>(&u);
#pragma unroll
                                for (int j = 0; j < 8; ++j)
                                    if (!(__bfloat162float(hm[j]) > 0.f)) z[q * 8 + j] = 0.f;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (i < nv && !(__bfloat162float(mk[i]) > 0.f)) z[i] = 0.f;
                        }
                    }
                    if (p.out_bf16) epi_store_bf16(p.out_bf16 + row * p.ld_out + col0, z, nv);
                    if (p.out_f32) {
                        if (p.atomic_out) {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (i < nv) atomicAdd(p.out_f32 + row * p.ld_out + col0 + i, z[i]);
                        } else {
                            epi_store_f32(p.out_f32 + row * p.ld_out + col0, z, nv);
                        }
                    }
                }
            }
        } else {
            // LayerNorm over the whole row (N == BN == 256, n0 == 0). Three TMEM passes; z parked in TMEM.
            float sum = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) z[i] = __uint_as_float(v[i]);
                if (row_ok) epi_bias_res(z, p, res_row, c * 32);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    sum += z[i];
                    v[i] = __float_as_uint(z[i]);
                }
                tmem_st_32x32b_x32(taddr + c * 32, v);
                if (row_ok && p.pre_ln) epi_store_bf16(p.pre_ln + row * p.ld_out + c * 32, z, 32);
            }
            tmem_st_wait();
            const float mean = sum * (1.f / BN);
            float var = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float d = __uint_as_float(v[i]) - mean;
                    var += d * d;
                }
            }
            const float rstd = rsqrtf(var * (1.f / BN) + p.ln_eps);
            if (row_ok) {
                if (p.ln_mean) p.ln_mean[row] = mean;
                if (p.ln_rstd) p.ln_rstd[row] = rstd;
            }
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                tmem_ld_32x32b_x32(taddr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    z[i] = (__uint_as_float(v[i]) - mean) * rstd * __ldg(p.ln_gamma + c * 32 + i) +
                           __ldg(p.ln_beta + c * 32 + i);
                if (row_ok) {
                    if (p.out_bf16) epi_store_bf16(p.out_bf16 + row * p.ld_out + c * 32, z, 32);
                    if (p.out_f32) epi_store_f32(p.out_f32 + row * p.ld_out + c * 32, z, 32);
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const fod_linear_args* a, GemmEpi epi, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
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
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(gemm_bf16_kernel<BN, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Cfg::kSmemBytes);
        attr_set = true;
    }
    const int total_kb = (a->K + GEMM_BK - 1) / GEMM_BK;
    int splits = a->split_k > 1 ? a->split_k : 1;
    if (splits > total_kb) splits = total_kb;
    epi.kb_per_split = (total_kb + splits - 1) / splits;
    splits = (total_kb + epi.kb_per_split - 1) / epi.kb_per_split;  // no empty split
    dim3 grid((a->N + BN - 1) / BN, (a->M + GEMM_BM - 1) / GEMM_BM, splits);
    gemm_bf16_kernel<BN, A_MN, B_MN><<<grid, GEMM_THREADS, Cfg::kSmemBytes, st>>>(tmA, tmB, epi);
    return fod_check_launch("gemm_bf16_kernel");
}

template <int BN>
static int launch_gemm_major(const fod_linear_args* a, const GemmEpi& e, cudaStream_t st) {
    if (a->a_mn && a->b_mn) return launch_gemm<BN, true, true>(a, e, st);
    if (a->b_mn) return launch_gemm<BN, false, true>(a, e, st);
    if (a->a_mn) return fod_fail(FOD_E_ARG, "gemm: A MN-major with B K-major is not instantiated");
    return launch_gemm<BN, false, false>(a, e, st);
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
    if (a->ln_gamma && (a->N != 256 || !a->ln_beta || a->relu || a->split_k > 1 || a->relu_mask))
        return fod_fail(FOD_E_ARG, "linear: LayerNorm epilogue needs N == 256, beta, no relu/mask/split");
    if (!a->out_bf16 && !a->out_f32) return fod_fail(FOD_E_ARG, "linear: no output");
    if (a->split_k > 1 && (!a->atomic_out || !a->out_f32 || a->out_bf16 || a->relu || a->bias || a->residual))
        return fod_fail(FOD_E_ARG, "linear: split_k needs atomic fp32 output and a plain epilogue");
    GemmEpi e;
    e.bias = a->bias;
    e.residual = reinterpret_cast<const __nv_bfloat16*>(a->residual);
    e.relu_mask = reinterpret_cast<const __nv_bfloat16*>(a->relu_mask);
    e.col_scale = a->col_scale;
    e.atomic_out = a->atomic_out;
    e.kb_per_split = 0;
    e.ln_gamma = a->ln_gamma;
    e.ln_beta = a->ln_beta;
    e.out_bf16 = reinterpret_cast<__nv_bfloat16*>(a->out_bf16);
    e.out_f32 = a->out_f32;
    e.pre_ln = reinterpret_cast<__nv_bfloat16*>(a->pre_ln_bf16);
    e.ln_mean = a->ln_mean;
    e.ln_rstd = a->ln_rstd;
    e.ld_res = a->ld_res;
    e.ld_out = a->ld_out;
    e.M = a->M;
    e.N = a->N;
    e.K = a->K;
    e.res_row_mod = a->res_row_mod;
    e.relu = a->relu;
    e.ln_eps = a->ln_eps;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (a->ln_gamma || a->N > 128) return launch_gemm_major<256>(a, e, st);
    if (a->N > 64) return launch_gemm_major<128>(a, e, st);
    return launch_gemm_major<64>(a, e, st);
}
