// Flash-attention backward for sm_100a (TMA + tcgen05/TMEM), head dim 32 per part, NPARTS = 2 for the
// ConditionalDETR concat form.  Two kernels, both recomputing S = Q·Kᵀ and dP = dO·Vᵀ on the tensor cores:
//   attn_bwd_dq_kernel   : CTA per (batch, head, 128-query tile), streams K/V tiles:  dQ_p += dS · K_p
//   attn_bwd_dkdv_kernel : CTA per (batch, head, 128-key tile),   streams Q/dO tiles: dV += Pᵀ·dO, dK_p += dSᵀ·Q_p
// with P = exp(scale·S − lse), dS = scale · P ⊙ (dP − delta), delta = rowsum(dO ⊙ O)  (attention.py:350-372 backward).
// No atomics: each gradient element is produced by exactly one CTA, accumulation lives in TMEM.
// Warp roles as in the forward kernel: warps 0-3 element-wise (one TMEM lane = one row per thread),
// warp 4 TMA producer, warp 5 TMEM allocator + single-thread MMA issuer.
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"
#include "dropout.cuh"

namespace fod {

constexpr int AB_T = 128;                       // tile edge (queries and keys)
constexpr int AB_D = 32;                        // head dim per part
constexpr int AB_THREADS = 320;  // 8 softmax warps (two per TMEM lane quarter, 64 of the 128 score columns each: one
                                 // warp per scheduler cannot hide the ex2 / TMEM-load latencies), TMA warp, MMA warp
constexpr int AB_PROD = 8, AB_MMA = 9;
constexpr int AB_TILE = AB_T * AB_D * 2;        // 8 KB operand tile, 64-byte rows, SWIZZLE_64B
constexpr int AB_PT = AB_T * AB_T * 2;          // 32 KB P / dS tile: two SWIZZLE_128B slabs [128][64]
constexpr int AB_STAGES = 2;
constexpr float AB_LOG2E = 1.4426950408889634f;

struct AttnBwdParams {
    const float* lse;    // (B,H,Lq)
    const float* delta;  // (B,H,Lq)
    __nv_bfloat16 *dq1, *dq2, *dk1, *dk2, *dv;  // dense (B, L, H*32)
    long long ld_dq, ld_dk;  // row strides of dq1 and of dk1/dv (dq2, dk2 dense)
    int B, H, Lq, Lk;
    int bm_q1, bm_k1, bm_q2, bm_k2, bm_v;
    float scale;
    DropoutCfg drop;  // state != nullptr: the forward call's dropout on P (mask index ((b*H+h)*Lq + q) * roundup8(Lk) + k)
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void store_row32_bf16(__nv_bfloat16* dst, const uint32_t (&v)[32]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(v[q * 8 + 0]), __uint_as_float(v[q * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(v[q * 8 + 2]), __uint_as_float(v[q * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(v[q * 8 + 4]), __uint_as_float(v[q * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(v[q * 8 + 6]), __uint_as_float(v[q * 8 + 7]));
        reinterpret_cast<uint4*>(dst)[q] = u;
    }
}

// write 32 consecutive bf16 of row `row`, logical columns [c*32, c*32+32), into a [128][128] tile stored as two
// SWIZZLE_128B slabs of 64 columns
__device__ __forceinline__ void store_tile_chunk(uint8_t* tile, int row, int c, const float (&x)[32]) {
    uint8_t* slab = tile + (c >> 1) * (AB_T * 128);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(x[q * 8 + 0], x[q * 8 + 1]);
        u.y = pack_bf16x2(x[q * 8 + 2], x[q * 8 + 3]);
        u.z = pack_bf16x2(x[q * 8 + 4], x[q * 8 + 5]);
        u.w = pack_bf16x2(x[q * 8 + 6], x[q * 8 + 7]);
        *reinterpret_cast<uint4*>(slab + swz_off<128>(row, (c & 1) * 4 + q)) = u;
    }
}

// ================================================================================================ dQ
template <int NPARTS>
struct DqSmem {
    static constexpr int kQ = 0;                                        // Q parts
    static constexpr int kDO = kQ + NPARTS * AB_TILE;                   // dO tile
    static constexpr int kK = kDO + AB_TILE;                            // K parts, staged
    static constexpr int kV = kK + AB_STAGES * NPARTS * AB_TILE;        // V, staged
    static constexpr int kDS = kV + AB_STAGES * AB_TILE;                // dS tile
    static constexpr int kBars = kDS + AB_PT;
    static constexpr int kTotal = kBars + 128 + 1024;
};

template <int NPARTS>
__global__ void __launch_bounds__(AB_THREADS, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQ1, const __grid_constant__ CUtensorMap tmK1,
                   const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmK2,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                   const AttnBwdParams p) {
    using L = DqSmem<NPARTS>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* q_full = bars + 0;
    uint64_t* kv_full = bars + 1;   // [2]
    uint64_t* kv_empty = bars + 3;  // [2]
    uint64_t* s_full = bars + 5;
    uint64_t* ds_full = bars + 6;
    uint64_t* dq_done = bars + 7;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * AB_T, h = blockIdx.y, b = blockIdx.z;
    const int nkt = (p.Lk + AB_T - 1) / AB_T;

    if (warp == AB_PROD && lane == 0) {
        tma_prefetch_desc(&tmQ1);
        tma_prefetch_desc(&tmK1);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmDO);
        mbar_init(q_full, 1);
        for (int s = 0; s < AB_STAGES; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(ds_full, 256);
        mbar_init(dq_done, 1);
        fence_barrier_init();
    }
    if (warp == AB_MMA) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();
    const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dQ = tmem_base + 256;

    if (warp == AB_PROD) {
        if (lane == 0) {
            mbar_arrive_expect_tx(q_full, (NPARTS + 1) * AB_TILE);
            tma_load_3d(smem + L::kQ, &tmQ1, q_full, h * AB_D, q0, b * p.bm_q1);
            if (NPARTS == 2) tma_load_3d(smem + L::kQ + AB_TILE, &tmQ2, q_full, h * AB_D, q0, b * p.bm_q2);
            tma_load_3d(smem + L::kDO, &tmDO, q_full, h * AB_D, q0, b);
            for (int j = 0; j < nkt; ++j) {
                const int s = j % AB_STAGES;
                mbar_wait(kv_empty + s, ((j / AB_STAGES) & 1) ^ 1);
                mbar_arrive_expect_tx(kv_full + s, (NPARTS + 1) * AB_TILE);
                uint8_t* sk = smem + L::kK + s * NPARTS * AB_TILE;
                tma_load_3d(sk, &tmK1, kv_full + s, h * AB_D, j * AB_T, b * p.bm_k1);
                if (NPARTS == 2) tma_load_3d(sk + AB_TILE, &tmK2, kv_full + s, h * AB_D, j * AB_T, b * p.bm_k2);
                tma_load_3d(smem + L::kV + s * AB_TILE, &tmV, kv_full + s, h * AB_D, j * AB_T, b * p.bm_v);
            }
        }
    } else if (warp == AB_MMA) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(AB_T, AB_T, 0, 0);
            constexpr uint32_t idesc_g = umma_idesc_bf16(AB_T, AB_D, 0, 1);  // B operand MN-major
            const uint32_t q_addr = smem_u32(smem + L::kQ);
            const uint32_t do_addr = smem_u32(smem + L::kDO);
            const uint32_t ds_addr = smem_u32(smem + L::kDS);
            auto issue_S = [&](int j) {
                const int s = j % AB_STAGES;
                mbar_wait(kv_full + s, (j / AB_STAGES) & 1);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(smem + L::kK + s * NPARTS * AB_TILE);
                const uint32_t v_addr = smem_u32(smem + L::kV + s * AB_TILE);
#pragma unroll
                for (int part = 0; part < NPARTS; ++part)
#pragma unroll
                    for (int k = 0; k < AB_D / 16; ++k)
                        umma_bf16_ss(tmem_S, umma_smem_desc(q_addr + part * AB_TILE + k * 32, 16, 512, 4),
                                     umma_smem_desc(k_addr + part * AB_TILE + k * 32, 16, 512, 4), idesc_s,
                                     (part | k) ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < AB_D / 16; ++k)
                    umma_bf16_ss(tmem_dP, umma_smem_desc(do_addr + k * 32, 16, 512, 4),
                                 umma_smem_desc(v_addr + k * 32, 16, 512, 4), idesc_s, k ? 1u : 0u);
                umma_commit(s_full);
            };
            mbar_wait(q_full, 0);
            issue_S(0);
            for (int j = 0; j < nkt; ++j) {
                const int s = j % AB_STAGES;
                mbar_wait(ds_full, j & 1);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(smem + L::kK + s * NPARTS * AB_TILE);
#pragma unroll
                for (int part = 0; part < NPARTS; ++part)
#pragma unroll
                    for (int kk = 0; kk < AB_T / 16; ++kk)
                        umma_bf16_ss(tmem_dQ + part * AB_D,
                                     umma_smem_desc(ds_addr + (kk >> 2) * (AB_T * 128) + (kk & 3) * 32, 16, 1024, 2),
                                     umma_smem_desc(k_addr + part * AB_TILE + kk * 1024, 4096, 512, 4), idesc_g,
                                     (j | kk) ? 1u : 0u);
                umma_commit(kv_empty + s);
                if (j + 1 < nkt) issue_S(j + 1);
            }
            umma_commit(dq_done);
        }
    } else {
        const int half = warp >> 2;  // which 64 of the 128 score columns
        const int row = (warp & 3) * 32 + lane;
        const uint32_t lane_off = uint32_t((warp & 3) * 32) << 16;
        const int q = q0 + row;
        const bool q_ok = q < p.Lq;
        const long long st = ((long long)b * p.H + h) * p.Lq + q;
        const float lse2 = q_ok ? p.lse[st] * AB_LOG2E : INFINITY;  // +inf -> P = 0 for padded query rows
        const float dlt = q_ok ? p.delta[st] : 0.f;
        const float sl2 = p.scale * AB_LOG2E;
        uint8_t* sDS = smem + L::kDS;
        uint32_t vs[32], vp[32];
        PhiloxKey key = {0u, 0u, 0u, 0u};
        if (p.drop.state) key = dropout_key(p.drop);
        const unsigned long long g_row =
            (((unsigned long long)b * p.H + h) * p.Lq + (unsigned long long)min(q, p.Lq - 1)) * (unsigned long long)((p.Lk + 7) >> 3);
        for (int j = 0; j < nkt; ++j) {
            mbar_wait(s_full, j & 1);  // also implies the dQ MMAs of tile j-1 (reading sDS) have retired
            tc_fence_after();
            const int kvalid = p.Lk - j * AB_T;
#pragma unroll 1
            for (int c = 2 * half; c < 2 * half + 2; ++c) {
                tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, vs);
                tmem_ld_32x32b_x32(tmem_dP + lane_off + c * 32, vp);
                tmem_ld_wait();
                float ds[32];
                if (p.drop.state) {  // dP = (dO·Vᵀ) ⊙ mask / (1-p)
#pragma unroll
                    for (int qd = 0; qd < 4; ++qd) {
                        const Keep8 keep = dropout_draw8(key, g_row + (unsigned long long)(j * (AB_T / 8) + c * 4 + qd), p.drop.thresh);
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            vp[qd * 8 + e] = __float_as_uint(keep[e] ? __uint_as_float(vp[qd * 8 + e]) * p.drop.scale : 0.f);
                    }
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float pr = fast_exp2(fmaf(__uint_as_float(vs[i]), sl2, -lse2));
                    const float d = pr * (__uint_as_float(vp[i]) - dlt) * p.scale;
                    ds[i] = (c * 32 + i < kvalid) ? d : 0.f;
                }
                store_tile_chunk(sDS, row, c, ds);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(ds_full);
        }
        mbar_wait(dq_done, 0);
        tc_fence_after();
#pragma unroll
        for (int part = 0; part < NPARTS; ++part) {
            if (half != part) continue;  // warp-uniform: part 0 by the first four warps, part 1 by the others
            tmem_ld_32x32b_x32(tmem_dQ + lane_off + part * AB_D, vs);
            tmem_ld_wait();
            if (q_ok) {
                const long long ldg = part ? (long long)(p.H * AB_D) : p.ld_dq;
                __nv_bfloat16* dst = (part ? p.dq2 : p.dq1) + ((long long)b * p.Lq + q) * ldg + h * AB_D;
                store_row32_bf16(dst, vs);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == AB_MMA) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// ================================================================================================ dK, dV
template <int NPARTS>
struct DkvSmem {
    static constexpr int kK = 0;                                          // K parts (resident)
    static constexpr int kV = kK + NPARTS * AB_TILE;                      // V (resident)
    static constexpr int kQ = kV + AB_TILE;                               // Q parts, staged
    static constexpr int kDO = kQ + AB_STAGES * NPARTS * AB_TILE;         // dO, staged
    static constexpr int kPT = kDO + AB_STAGES * AB_TILE;                 // Pᵀ tile
    static constexpr int kDST = kPT + AB_PT;                              // dSᵀ tile
    static constexpr int kStat = kDST + AB_PT;                            // lse / delta rows, double buffered
    static constexpr int kBars = kStat + 2 * 2 * AB_T * 4;
    static constexpr int kTotal = kBars + 128 + 1024;
};

template <int NPARTS>
__global__ void __launch_bounds__(AB_THREADS, 1)
attn_bwd_dkdv_kernel(const __grid_constant__ CUtensorMap tmQ1, const __grid_constant__ CUtensorMap tmK1,
                     const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmK2,
                     const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                     const AttnBwdParams p) {
    using L = DkvSmem<NPARTS>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* kv_full = bars + 0;
    uint64_t* q_full = bars + 1;   // [2]
    uint64_t* q_empty = bars + 3;  // [2]
    uint64_t* s_full = bars + 5;
    uint64_t* pds_full = bars + 6;
    uint64_t* done = bars + 7;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    float* stat = reinterpret_cast<float*>(smem + L::kStat);  // [buf][lse|delta][128]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k0 = blockIdx.x * AB_T, h = blockIdx.y, b = blockIdx.z;
    const int nqt = (p.Lq + AB_T - 1) / AB_T;

    if (warp == AB_PROD && lane == 0) {
        tma_prefetch_desc(&tmQ1);
        tma_prefetch_desc(&tmK1);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmDO);
        mbar_init(kv_full, 1);
        for (int s = 0; s < AB_STAGES; ++s) {
            mbar_init(q_full + s, 1);
            mbar_init(q_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(pds_full, 256);
        mbar_init(done, 1);
        fence_barrier_init();
    }
    if (warp == AB_MMA) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();
    const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dV = tmem_base + 256, tmem_dK = tmem_base + 288;

    if (warp == AB_PROD) {
        if (lane == 0) {
            mbar_arrive_expect_tx(kv_full, (NPARTS + 1) * AB_TILE);
            tma_load_3d(smem + L::kK, &tmK1, kv_full, h * AB_D, k0, b * p.bm_k1);
            if (NPARTS == 2) tma_load_3d(smem + L::kK + AB_TILE, &tmK2, kv_full, h * AB_D, k0, b * p.bm_k2);
            tma_load_3d(smem + L::kV, &tmV, kv_full, h * AB_D, k0, b * p.bm_v);
            for (int i = 0; i < nqt; ++i) {
                const int s = i % AB_STAGES;
                mbar_wait(q_empty + s, ((i / AB_STAGES) & 1) ^ 1);
                mbar_arrive_expect_tx(q_full + s, (NPARTS + 1) * AB_TILE);
                uint8_t* sq = smem + L::kQ + s * NPARTS * AB_TILE;
                tma_load_3d(sq, &tmQ1, q_full + s, h * AB_D, i * AB_T, b * p.bm_q1);
                if (NPARTS == 2) tma_load_3d(sq + AB_TILE, &tmQ2, q_full + s, h * AB_D, i * AB_T, b * p.bm_q2);
                tma_load_3d(smem + L::kDO + s * AB_TILE, &tmDO, q_full + s, h * AB_D, i * AB_T, b);
            }
        }
    } else if (warp == AB_MMA) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(AB_T, AB_T, 0, 0);
            constexpr uint32_t idesc_g = umma_idesc_bf16(AB_T, AB_D, 0, 1);
            const uint32_t k_addr = smem_u32(smem + L::kK);
            const uint32_t v_addr = smem_u32(smem + L::kV);
            const uint32_t pt_addr = smem_u32(smem + L::kPT);
            const uint32_t dst_addr = smem_u32(smem + L::kDST);
            auto issue_S = [&](int i) {  // Sᵀ = K·Qᵀ, dPᵀ = V·dOᵀ
                const int s = i % AB_STAGES;
                mbar_wait(q_full + s, (i / AB_STAGES) & 1);
                tc_fence_after();
                const uint32_t q_addr = smem_u32(smem + L::kQ + s * NPARTS * AB_TILE);
                const uint32_t do_addr = smem_u32(smem + L::kDO + s * AB_TILE);
#pragma unroll
                for (int part = 0; part < NPARTS; ++part)
#pragma unroll
                    for (int k = 0; k < AB_D / 16; ++k)
                        umma_bf16_ss(tmem_S, umma_smem_desc(k_addr + part * AB_TILE + k * 32, 16, 512, 4),
                                     umma_smem_desc(q_addr + part * AB_TILE + k * 32, 16, 512, 4), idesc_s,
                                     (part | k) ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < AB_D / 16; ++k)
                    umma_bf16_ss(tmem_dP, umma_smem_desc(v_addr + k * 32, 16, 512, 4),
                                 umma_smem_desc(do_addr + k * 32, 16, 512, 4), idesc_s, k ? 1u : 0u);
                umma_commit(s_full);
            };
            mbar_wait(kv_full, 0);
            issue_S(0);
            for (int i = 0; i < nqt; ++i) {
                const int s = i % AB_STAGES;
                mbar_wait(pds_full, i & 1);
                tc_fence_after();
                const uint32_t q_addr = smem_u32(smem + L::kQ + s * NPARTS * AB_TILE);
                const uint32_t do_addr = smem_u32(smem + L::kDO + s * AB_TILE);
#pragma unroll
                for (int kk = 0; kk < AB_T / 16; ++kk)
                    umma_bf16_ss(tmem_dV, umma_smem_desc(pt_addr + (kk >> 2) * (AB_T * 128) + (kk & 3) * 32, 16, 1024, 2),
                                 umma_smem_desc(do_addr + kk * 1024, 4096, 512, 4), idesc_g, (i | kk) ? 1u : 0u);
#pragma unroll
                for (int part = 0; part < NPARTS; ++part)
#pragma unroll
                    for (int kk = 0; kk < AB_T / 16; ++kk)
                        umma_bf16_ss(tmem_dK + part * AB_D,
                                     umma_smem_desc(dst_addr + (kk >> 2) * (AB_T * 128) + (kk & 3) * 32, 16, 1024, 2),
                                     umma_smem_desc(q_addr + part * AB_TILE + kk * 1024, 4096, 512, 4), idesc_g,
                                     (i | kk) ? 1u : 0u);
                umma_commit(q_empty + s);
                if (i + 1 < nqt) issue_S(i + 1);
            }
            umma_commit(done);
        }
    } else {
        const int half = warp >> 2;        // which 64 of the 128 query columns of Sᵀ / dPᵀ
        const int row = (warp & 3) * 32 + lane;  // key row
        const uint32_t lane_off = uint32_t((warp & 3) * 32) << 16;
        const float sl2 = p.scale * AB_LOG2E;
        uint8_t* sPT = smem + L::kPT;
        uint8_t* sDST = smem + L::kDST;
        uint32_t vs[32], vp[32];
        const long long st0 = ((long long)b * p.H + h) * p.Lq;
        PhiloxKey key = {0u, 0u, 0u, 0u};
        if (p.drop.state) key = dropout_key(p.drop);
        const unsigned long long lkp8 = (unsigned long long)((p.Lk + 7) >> 3);
        const unsigned long long kgrp = (unsigned long long)((k0 + row) >> 3);  // this key's group of 8 within a score row
        for (int i = 0; i < nqt; ++i) {
            float* lse_s = stat + (i & 1) * 2 * AB_T;
            float* dl_s = lse_s + AB_T;
            if (half == 0) {
                const int q = i * AB_T + row;
                const bool ok = q < p.Lq;
                lse_s[row] = ok ? p.lse[st0 + q] * AB_LOG2E : INFINITY;  // padded query columns -> P = 0
                dl_s[row] = ok ? p.delta[st0 + q] : 0.f;
            }
            named_bar_sync(1, 256);
            mbar_wait(s_full, i & 1);  // also implies tile i-1's dV/dK MMAs (reading sPT/sDST) have retired
            tc_fence_after();
#pragma unroll 1
            for (int c = 2 * half; c < 2 * half + 2; ++c) {
                tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, vs);
                tmem_ld_32x32b_x32(tmem_dP + lane_off + c * 32, vp);
                tmem_ld_wait();
                float pt[32], ds[32];
                uint32_t keepbits = 0xffffffffu;  // bit j: the weight of (query c*32 + j, this key) was kept
                if (p.drop.state) {
                    // The mask groups run along the keys: the 8 lanes that share a key group each draw the group of ONE of
                    // 8 consecutive queries and hand the byte round by shuffle — one Philox call per 8 elements, as in
                    // the forward kernel.
                    keepbits = 0u;
#pragma unroll
                    for (int q8 = 0; q8 < 4; ++q8) {
                        const int qq = min(i * AB_T + c * 32 + q8 * 8 + (lane & 7), p.Lq - 1);
                        const uint32_t mine = dropout_keep8(key, ((unsigned long long)st0 + qq) * lkp8 + kgrp, p.drop.thresh);
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            const uint32_t m = __shfl_sync(0xffffffffu, mine, (lane & ~7) | t);
                            keepbits |= ((m >> (row & 7)) & 1u) << (q8 * 8 + t);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float mk = p.drop.state ? (((keepbits >> j) & 1u) ? p.drop.scale : 0.f) : 1.f;
                    const float pr = fast_exp2(fmaf(__uint_as_float(vs[j]), sl2, -lse_s[c * 32 + j]));
                    pt[j] = pr * mk;
                    ds[j] = pr * (__uint_as_float(vp[j]) * mk - dl_s[c * 32 + j]) * p.scale;
                }
                store_tile_chunk(sPT, row, c, pt);
                store_tile_chunk(sDST, row, c, ds);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(pds_full);
        }
        mbar_wait(done, 0);
        tc_fence_after();
        const int key_row = k0 + row;
        const bool k_ok = key_row < p.Lk;
        const long long orow = ((long long)b * p.Lk + key_row) * p.ld_dk + h * AB_D;
        const long long orow2 = ((long long)b * p.Lk + key_row) * (p.H * AB_D) + h * AB_D;
        if (half == 0) {  // warp-uniform split of the read-out: dV (and dK part 1) by warps 0-3, dK part 0 by warps 4-7
            tmem_ld_32x32b_x32(tmem_dV + lane_off, vs);
            tmem_ld_wait();
            if (k_ok) store_row32_bf16(p.dv + orow, vs);
        }
#pragma unroll
        for (int part = 0; part < NPARTS; ++part) {
            if (half == part) continue;
            tmem_ld_32x32b_x32(tmem_dK + lane_off + part * AB_D, vs);
            tmem_ld_wait();
            if (k_ok) store_row32_bf16(part ? p.dk2 + orow2 : p.dk1 + orow, vs);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == AB_MMA) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// delta[b,h,q] = sum_d dO[b,q,h*32+d] * O[b,q,h*32+d]   (one thread per (b,q,h); 64-byte segments)
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                                  long long ldo, long long bso, long long lddo, long long bsdo,
                                  float* __restrict__ delta, int B, int H, int Lq) {
    pdl_launch_dependents();
    pdl_wait();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)B * Lq * H;
    if (idx >= total) return;
    const int h = (int)(idx % H);
    const long long bq = idx / H;
    const int q = (int)(bq % Lq), b = (int)(bq / Lq);
    const uint4* po = reinterpret_cast<const uint4*>(o + b * bso + (long long)q * ldo + h * AB_D);
    const uint4* pd = reinterpret_cast<const uint4*>(dout + b * bsdo + (long long)q * lddo + h * AB_D);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint4 a = __ldg(po + i), c = __ldg(pd + i);
        const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
        const __nv_bfloat162* hc = reinterpret_cast<const __nv_bfloat162*>(&c);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 x = __bfloat1622float2(ha[j]), y = __bfloat1622float2(hc[j]);
            s = fmaf(x.x, y.x, s);
            s = fmaf(x.y, y.y, s);
        }
    }
    delta[((long long)b * H + h) * Lq + q] = s;
}

static int make_seq_tmap_b(CUtensorMap* tm, const void* base, int rowlen, int L, int B, long long ld, long long bs) {
    const bool bcast = (bs == 0) || (B == 1);
    uint64_t dims[3] = {(uint64_t)rowlen, (uint64_t)L, (uint64_t)(bcast ? 1 : B)};
    uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)(bcast ? (uint64_t)ld * 2 * L : (uint64_t)bs * 2)};
    uint32_t box[3] = {AB_D, AB_T, 1};
    return make_tmap_bf16(tm, base, 3, dims, str, box, 64);
}

}  // namespace fod

extern "C" int fod_attn_bwd(const fod_attn_bwd_args* g, void* stream) {
    using namespace fod;
    if (!g) return fod_fail(FOD_E_ARG, "attn_bwd: null args");
    const fod_attn_args* a = &g->fwd;
    if (!a->q1 || !a->k1 || !a->v || !a->out || !a->lse || !g->dout || !g->delta || !g->dq1 || !g->dk1 || !g->dv)
        return fod_fail(FOD_E_ARG, "attn_bwd: null pointer");
    if (a->B <= 0 || a->H <= 0 || a->Lq <= 0 || a->Lk <= 0) return fod_fail(FOD_E_ARG, "attn_bwd: empty shape");
    const int two = a->q2 != nullptr;
    if (two && (!a->k2 || !g->dq2 || !g->dk2)) return fod_fail(FOD_E_ARG, "attn_bwd: second part pointers");
    const long long chk[] = {a->ldq1, a->ldk1, a->ldv, a->ldo, g->lddo, a->bsq1, a->bsk1, a->bsv, a->bso, g->bsdo,
                             two ? a->ldq2 : 8, two ? a->ldk2 : 8, two ? a->bsq2 : 0, two ? a->bsk2 : 0};
    for (long long x : chk)
        if (x % 8) return fod_fail(FOD_E_ARG, "attn_bwd: strides must be multiples of 8 elements");
    {
        // one fused kernel (attn_bwd_fused_sm100.cu) wherever dQ of all query tiles fits TMEM; FOD_ATTN_BWD_FUSED=0
        // (or mode 1) keeps the two-kernel path
        const long long gchk[] = {g->ld_dq, g->ld_dk};
        for (long long x : gchk)
            if (x % 8) return fod_fail(FOD_E_ARG, "attn_bwd: gradient strides must be multiples of 8");
        int mode = g->mode;
        if (mode < 0 || mode > 2) return fod_fail(FOD_E_ARG, "attn_bwd: mode must be 0, 1 or 2");
        if (mode == 2 && !attn_bwd_fused_supported(g)) return fod_fail(FOD_E_ARG, "attn_bwd: sequence too long for the fused kernel");
        if (mode == 0) mode = (fod_env_int("FOD_ATTN_BWD_FUSED", 1) && attn_bwd_fused_supported(g)) ? 2 : 1;
        if (mode == 2) return launch_attn_bwd_fused(g, reinterpret_cast<cudaStream_t>(stream));
    }
    const int rowlen = a->H * AB_D;
    CUtensorMap tq1, tk1, tq2, tk2, tv, tdo;
    if (make_seq_tmap_b(&tq1, a->q1, rowlen, a->Lq, a->B, a->ldq1, a->bsq1)) return fod_fail(FOD_E_TMAP, "attn_bwd: q1");
    if (make_seq_tmap_b(&tk1, a->k1, rowlen, a->Lk, a->B, a->ldk1, a->bsk1)) return fod_fail(FOD_E_TMAP, "attn_bwd: k1");
    if (make_seq_tmap_b(&tv, a->v, rowlen, a->Lk, a->B, a->ldv, a->bsv)) return fod_fail(FOD_E_TMAP, "attn_bwd: v");
    if (make_seq_tmap_b(&tdo, g->dout, rowlen, a->Lq, a->B, g->lddo, a->B > 1 ? g->bsdo : 0))
        return fod_fail(FOD_E_TMAP, "attn_bwd: dout");
    if (two) {
        if (make_seq_tmap_b(&tq2, a->q2, rowlen, a->Lq, a->B, a->ldq2, a->bsq2)) return fod_fail(FOD_E_TMAP, "attn_bwd: q2");
        if (make_seq_tmap_b(&tk2, a->k2, rowlen, a->Lk, a->B, a->ldk2, a->bsk2)) return fod_fail(FOD_E_TMAP, "attn_bwd: k2");
    } else {
        tq2 = tq1;
        tk2 = tk1;
    }
    AttnBwdParams p;
    p.lse = a->lse;
    p.delta = g->delta;
    p.dq1 = reinterpret_cast<__nv_bfloat16*>(g->dq1);
    p.dq2 = reinterpret_cast<__nv_bfloat16*>(g->dq2);
    p.dk1 = reinterpret_cast<__nv_bfloat16*>(g->dk1);
    p.dk2 = reinterpret_cast<__nv_bfloat16*>(g->dk2);
    p.dv = reinterpret_cast<__nv_bfloat16*>(g->dv);
    p.B = a->B; p.H = a->H; p.Lq = a->Lq; p.Lk = a->Lk;
    p.ld_dq = g->ld_dq ? g->ld_dq : rowlen;
    p.ld_dk = g->ld_dk ? g->ld_dk : rowlen;
    if ((p.ld_dq % 8) || (p.ld_dk % 8)) return fod_fail(FOD_E_ARG, "attn_bwd: gradient strides must be multiples of 8");
    auto has_b = [&](long long bs) { return (bs != 0 && a->B > 1) ? 1 : 0; };
    p.bm_q1 = has_b(a->bsq1); p.bm_k1 = has_b(a->bsk1);
    p.bm_q2 = two ? has_b(a->bsq2) : 0; p.bm_k2 = two ? has_b(a->bsk2) : 0;
    p.bm_v = has_b(a->bsv);
    p.scale = a->scale;
    p.drop = {nullptr, 0u, 0u, 1.f};
    if (a->drop.rng_state && a->drop.p > 0.f) {
        if (!(a->drop.p < 1.f)) return fod_fail(FOD_E_ARG, "attn_bwd: dropout p must be in [0, 1)");
        p.drop.state = reinterpret_cast<const unsigned long long*>(a->drop.rng_state);
        p.drop.call = (unsigned int)a->drop.call;
        p.drop.thresh = (unsigned int)(a->drop.p * 65536.f + 0.5f);
        p.drop.scale = 1.f / (1.f - a->drop.p);
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    {
        const long long total = (long long)a->B * a->Lq * a->H;
        fod_launch(attn_delta_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st,
            reinterpret_cast<const __nv_bfloat16*>(a->out), reinterpret_cast<const __nv_bfloat16*>(g->dout), a->ldo,
            a->bso, g->lddo, g->bsdo, g->delta, a->B, a->H, a->Lq);
        int rc = fod_check_launch("attn_delta_kernel");
        if (rc) return rc;
    }
    static bool attr[2] = {false, false};
    if (!attr[two]) {
        if (two) {
            cudaFuncSetAttribute(attn_bwd_dq_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DqSmem<2>::kTotal);
            cudaFuncSetAttribute(attn_bwd_dkdv_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DkvSmem<2>::kTotal);
        } else {
            cudaFuncSetAttribute(attn_bwd_dq_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, DqSmem<1>::kTotal);
            cudaFuncSetAttribute(attn_bwd_dkdv_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, DkvSmem<1>::kTotal);
        }
        attr[two] = true;
    }
    dim3 gq((a->Lq + AB_T - 1) / AB_T, a->H, a->B), gk((a->Lk + AB_T - 1) / AB_T, a->H, a->B);
    if (two) {
        fod_launch(attn_bwd_dq_kernel<2>, gq, dim3(AB_THREADS), DqSmem<2>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
        fod_launch(attn_bwd_dkdv_kernel<2>, gk, dim3(AB_THREADS), DkvSmem<2>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
    } else {
        fod_launch(attn_bwd_dq_kernel<1>, gq, dim3(AB_THREADS), DqSmem<1>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
        fod_launch(attn_bwd_dkdv_kernel<1>, gk, dim3(AB_THREADS), DkvSmem<1>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
    }
    return fod_check_launch("attn_bwd kernels");
}
