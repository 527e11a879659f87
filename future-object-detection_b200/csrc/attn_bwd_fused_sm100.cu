// Fused flash-attention backward for sm_100a: ONE kernel per call, one recompute of S and dP per (query tile, key tile).
//
//   CTA per (batch, head).  Outer loop over 128-key tiles j, inner loop over 128-query tiles i:
//     S = Q_i·K_jᵀ, dP = dO_i·V_jᵀ                      (tcgen05.mma, accumulators in TMEM, single buffer each)
//     P = exp(scale·S − lse), dS = scale·P ⊙ (dP − delta)  (8 softmax warps, one query row per thread, 64 columns each;
//                                                           both tiles are pulled into registers first, so the NEXT
//                                                           S / dP pair is issued while this one is exponentiated)
//     dV_j += Pᵀ·dO_i, dK_j += dSᵀ·Q_i, dQ_i += dS·K_j     (tcgen05.mma; P and dS are written ONCE to shared memory
//                                                           as [query][key] tiles and serve as K-major A operand for
//                                                           dQ and as MN-major A operand for dV / dK)
//   dQ of EVERY query tile stays resident in TMEM for the whole CTA (32 columns per tile and part), dK / dV of the
//   current key tile likewise; they are read out once.  No atomics, no second pass over S: 10·N·S²·D flop executed.
//   delta = rowsum(dO ⊙ O) is computed by the softmax warps during the first key tile (no separate kernel).
//
// TMEM: S 128 + dP 128 + dV 32 + dK 32·NP + dQ 32·NP·nqt  <= 512 columns  =>  nqt <= 6 (one part), nqt <= 2 (two parts).
// Longer sequences take the two-kernel path of attn_bwd_sm100.cu.
// Bound: MUFU ex2 (one exponential per 160 tensor flop at head dim 32): ~1024 clk per 128x128 tile pair and SM.
// Replaces the autograd of ConditionalDETR/models/attention.py:350-372 and of torch.nn.MultiheadAttention's core
// (future_od/models/transformer.py:404,417).
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"

namespace fod {

constexpr int AF_T = 128;
constexpr int AF_D = 32;
constexpr int AF_THREADS = 320;
constexpr int AF_PROD = 8, AF_MMA = 9;
constexpr int AF_TILE = AF_T * AF_D * 2;  // 8 KB operand tile, 64-byte rows, SWIZZLE_64B
constexpr int AF_PT = AF_T * AF_T * 2;    // 32 KB P / dS tile: two SWIZZLE_128B slabs [128 query][64 key]
constexpr float AF_LOG2E = 1.4426950408889634f;

template <int NP>
struct FusedBwdSmem {
    static constexpr int kMaxQT = NP == 1 ? 6 : 2;
    static constexpr int kK = 0;                            // K parts, 2 stages
    static constexpr int kV = kK + 2 * NP * AF_TILE;        // V, 2 stages
    static constexpr int kQ = kV + 2 * AF_TILE;             // Q parts, 2 stages
    static constexpr int kDO = kQ + 2 * NP * AF_TILE;       // dO, 2 stages
    static constexpr int kP = kDO + 2 * AF_TILE;            // P tile
    static constexpr int kDS = kP + AF_PT;                  // dS tile
    static constexpr int kStat = kDS + AF_PT;               // [2 halves][lse2 | delta][kMaxQT*128] fp32
    static constexpr int kBars = kStat + 2 * 2 * kMaxQT * AF_T * 4;
    static constexpr int kTotal = kBars + 256 + 1024;       // + alignment slack
};

struct AttnBwdFusedParams {
    const float* lse;            // (B,H,Lq)
    float* delta;                // (B,H,Lq) optional output (kept for API parity with the two-kernel path)
    const __nv_bfloat16* out;    // O (B,Lq,H*32)
    long long ldo, bso;
    __nv_bfloat16 *dq1, *dq2, *dk1, *dk2, *dv;
    long long ld_dq, ld_dk;
    int B, H, Lq, Lk;
    int bm_q1, bm_k1, bm_q2, bm_k2, bm_v;
    float scale;
};

__device__ __forceinline__ void af_store_row32(__nv_bfloat16* dst, const uint32_t (&v)[32]) {
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

template <int NP>
__global__ void __launch_bounds__(AF_THREADS, 1)
attn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tmQ1, const __grid_constant__ CUtensorMap tmK1,
                      const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmK2,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                      const AttnBwdFusedParams p) {
    using L = FusedBwdSmem<NP>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* kv_full = bars + 0;    // [2]
    uint64_t* kv_empty = bars + 2;   // [2]
    uint64_t* qdo_full = bars + 4;   // [2]
    uint64_t* qdo_empty = bars + 6;  // [2]
    uint64_t* s_full = bars + 8;     // S / dP of tile t are complete in TMEM
    uint64_t* s_free = bars + 9;     // ... and have been copied into registers by all 256 softmax threads
    uint64_t* p_full = bars + 10;    // P / dS of tile t are in shared memory
    uint64_t* g_done = bars + 11;    // the gradient MMAs of tile t have retired (P / dS / operand stages reusable)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    float* stat = reinterpret_cast<float*>(smem + L::kStat);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    const int nqt = (p.Lq + AF_T - 1) / AF_T;
    const int nkt = (p.Lk + AF_T - 1) / AF_T;
    const int ntiles = nqt * nkt;

    if (warp == AF_PROD && lane == 0) {
        tma_prefetch_desc(&tmQ1);
        tma_prefetch_desc(&tmK1);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmDO);
        if (NP == 2) {
            tma_prefetch_desc(&tmQ2);
            tma_prefetch_desc(&tmK2);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, 1);
            mbar_init(qdo_full + s, 1);
            mbar_init(qdo_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(s_free, 256);
        mbar_init(p_full, 256);
        mbar_init(g_done, 1);
        fence_barrier_init();
    }
    if (warp == AF_MMA) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();
    const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dV = tmem_base + 256, tmem_dK = tmem_base + 288;
    const uint32_t tmem_dQ = tmem_dK + NP * AF_D;  // + (i * NP + part) * 32

    if (warp == AF_PROD) {
        if (lane == 0) {
            int t = 0;
            for (int j = 0; j < nkt; ++j) {
                const int sj = j & 1;
                mbar_wait(kv_empty + sj, ((j >> 1) & 1) ^ 1);
                mbar_arrive_expect_tx(kv_full + sj, (NP + 1) * AF_TILE);
                uint8_t* sk = smem + L::kK + sj * NP * AF_TILE;
                tma_load_3d(sk, &tmK1, kv_full + sj, h * AF_D, j * AF_T, b * p.bm_k1);
                if (NP == 2) tma_load_3d(sk + AF_TILE, &tmK2, kv_full + sj, h * AF_D, j * AF_T, b * p.bm_k2);
                tma_load_3d(smem + L::kV + sj * AF_TILE, &tmV, kv_full + sj, h * AF_D, j * AF_T, b * p.bm_v);
                for (int i = 0; i < nqt; ++i, ++t) {
                    const int s = t & 1;
                    mbar_wait(qdo_empty + s, ((t >> 1) & 1) ^ 1);
                    mbar_arrive_expect_tx(qdo_full + s, (NP + 1) * AF_TILE);
                    uint8_t* sq = smem + L::kQ + s * NP * AF_TILE;
                    tma_load_3d(sq, &tmQ1, qdo_full + s, h * AF_D, i * AF_T, b * p.bm_q1);
                    if (NP == 2) tma_load_3d(sq + AF_TILE, &tmQ2, qdo_full + s, h * AF_D, i * AF_T, b * p.bm_q2);
                    tma_load_3d(smem + L::kDO + s * AF_TILE, &tmDO, qdo_full + s, h * AF_D, i * AF_T, b);
                }
            }
        }
    } else if (warp == AF_MMA) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(AF_T, AF_T, 0, 0);   // S, dP: both operands K-major
            constexpr uint32_t idesc_q = umma_idesc_bf16(AF_T, AF_D, 0, 1);   // dQ: A = dS K-major, B = K_j MN-major
            constexpr uint32_t idesc_k = umma_idesc_bf16(AF_T, AF_D, 1, 1);   // dV, dK: A = Pᵀ / dSᵀ (MN-major view)
            const uint32_t p_addr = smem_u32(smem + L::kP);
            const uint32_t ds_addr = smem_u32(smem + L::kDS);
            auto issue_SP = [&](int t) {
                const int j = t / nqt, i = t - j * nqt;
                const int sj = j & 1, s = t & 1;
                if (i == 0) mbar_wait(kv_full + sj, (j >> 1) & 1);
                mbar_wait(qdo_full + s, (t >> 1) & 1);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(smem + L::kK + sj * NP * AF_TILE);
                const uint32_t v_addr = smem_u32(smem + L::kV + sj * AF_TILE);
                const uint32_t q_addr = smem_u32(smem + L::kQ + s * NP * AF_TILE);
                const uint32_t do_addr = smem_u32(smem + L::kDO + s * AF_TILE);
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int k = 0; k < AF_D / 16; ++k)
                        umma_bf16_ss(tmem_S, umma_smem_desc(q_addr + part * AF_TILE + k * 32, 16, 512, 4),
                                     umma_smem_desc(k_addr + part * AF_TILE + k * 32, 16, 512, 4), idesc_s,
                                     (part | k) ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < AF_D / 16; ++k)
                    umma_bf16_ss(tmem_dP, umma_smem_desc(do_addr + k * 32, 16, 512, 4),
                                 umma_smem_desc(v_addr + k * 32, 16, 512, 4), idesc_s, k ? 1u : 0u);
                umma_commit(s_full);
            };
            issue_SP(0);
            for (int t = 0; t < ntiles; ++t) {
                const int j = t / nqt, i = t - j * nqt;
                const int sj = j & 1, s = t & 1;
                if (t + 1 < ntiles) {
                    // S(t) / dP(t) live in the softmax warps' registers: the next pair is computed while they work
                    mbar_wait(s_free, t & 1);
                    tc_fence_after();
                    issue_SP(t + 1);
                }
                mbar_wait(p_full, t & 1);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(smem + L::kK + sj * NP * AF_TILE);
                const uint32_t q_addr = smem_u32(smem + L::kQ + s * NP * AF_TILE);
                const uint32_t do_addr = smem_u32(smem + L::kDO + s * AF_TILE);
                // dV_j (+)= Pᵀ·dO_i: contraction over the 128 queries, 16 per instruction.  A: rows of the [query][key]
                // tile are the contraction index, 64 keys contiguous per slab (MN-major, slabs 16 KB apart).
#pragma unroll
                for (int kk = 0; kk < AF_T / 16; ++kk)
                    umma_bf16_ss(tmem_dV, umma_smem_desc(p_addr + kk * 2048, AF_T * 128, 1024, 2),
                                 umma_smem_desc(do_addr + kk * 1024, 4096, 512, 4), idesc_k, (i | kk) ? 1u : 0u);
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int kk = 0; kk < AF_T / 16; ++kk)
                        umma_bf16_ss(tmem_dK + part * AF_D, umma_smem_desc(ds_addr + kk * 2048, AF_T * 128, 1024, 2),
                                     umma_smem_desc(q_addr + part * AF_TILE + kk * 1024, 4096, 512, 4), idesc_k,
                                     (i | kk) ? 1u : 0u);
                // dQ_i (+)= dS·K_j: contraction over the 128 keys
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int kk = 0; kk < AF_T / 16; ++kk)
                        umma_bf16_ss(tmem_dQ + (i * NP + part) * AF_D,
                                     umma_smem_desc(ds_addr + (kk >> 2) * (AF_T * 128) + (kk & 3) * 32, 16, 1024, 2),
                                     umma_smem_desc(k_addr + part * AF_TILE + kk * 1024, 4096, 512, 4), idesc_q,
                                     (j | kk) ? 1u : 0u);
                umma_commit(g_done);
                umma_commit(qdo_empty + s);
                if (i == nqt - 1) umma_commit(kv_empty + sj);
            }
        }
    } else {
        // ------------------------------------------------------------------------------------------ softmax warps
        const int quarter = warp & 3, half = warp >> 2;  // TMEM lane quarter; which 64 of the 128 key columns
        const int row = quarter * 32 + lane;
        const uint32_t lane_off = uint32_t(quarter * 32) << 16;
        const float sl2 = p.scale * AF_LOG2E;
        float* my_lse = stat + (half * 2 + 0) * (L::kMaxQT * AF_T);
        float* my_dlt = stat + (half * 2 + 1) * (L::kMaxQT * AF_T);
        uint8_t* sP = smem + L::kP;
        uint8_t* sDS = smem + L::kDS;
        const long long C = (long long)p.H * AF_D;

        auto read_out_dkv = [&](int jj) {  // dV / dK of key tile jj: TMEM -> bf16 rows
            const int key = jj * AF_T + row;
            const bool k_ok = key < p.Lk;
            const long long orow = ((long long)b * p.Lk + key) * p.ld_dk + h * AF_D;
            const long long orow2 = ((long long)b * p.Lk + key) * C + h * AF_D;
            uint32_t v[32];
            if (half == 0) {
                tmem_ld_32x32b_x32(tmem_dV + lane_off, v);
                tmem_ld_wait();
                if (k_ok) af_store_row32(p.dv + orow, v);
                if (NP == 2) {
                    tmem_ld_32x32b_x32(tmem_dK + lane_off + AF_D, v);
                    tmem_ld_wait();
                    if (k_ok) af_store_row32(p.dk2 + orow2, v);
                }
            } else {
                tmem_ld_32x32b_x32(tmem_dK + lane_off, v);
                tmem_ld_wait();
                if (k_ok) af_store_row32(p.dk1 + orow, v);
            }
        };

        uint32_t vs[2][32], vp[2][32];
        for (int t = 0; t < ntiles; ++t) {
            const int j = t / nqt, i = t - j * nqt;
            const int q = i * AF_T + row;
            const bool q_ok = q < p.Lq;
            mbar_wait(s_full, t & 1);
            tc_fence_after();
            tmem_ld_32x32b_x32(tmem_S + lane_off + (2 * half) * 32, vs[0]);
            tmem_ld_32x32b_x32(tmem_S + lane_off + (2 * half + 1) * 32, vs[1]);
            tmem_ld_32x32b_x32(tmem_dP + lane_off + (2 * half) * 32, vp[0]);
            tmem_ld_32x32b_x32(tmem_dP + lane_off + (2 * half + 1) * 32, vp[1]);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(s_free);

            float lse2, dlt;
            if (j == 0) {
                // first key tile: this row's lse and delta = sum_d dO[q][d]·O[q][d]; both halves compute their own copy
                // (no cross-warp hand-over).  dO comes from the staged tile, O from global memory.
                const int s = t & 1;
                mbar_wait(qdo_full + s, (t >> 1) & 1);  // completed long ago; makes the TMA writes visible to this thread
                float acc = 0.f;
                if (q_ok) {
                    const uint8_t* sdo = smem + L::kDO + s * AF_TILE;
                    const uint4* po = reinterpret_cast<const uint4*>(p.out + (long long)b * p.bso + (long long)q * p.ldo + h * AF_D);
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const uint4 a = __ldg(po + c);
                        const uint4 d = *reinterpret_cast<const uint4*>(sdo + swz_off<64>(row, c));
                        const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
                        const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 x = __bfloat1622float2(ha[e]), y = __bfloat1622float2(hd[e]);
                            acc = fmaf(x.x, y.x, acc);
                            acc = fmaf(x.y, y.y, acc);
                        }
                    }
                }
                const long long st = ((long long)b * p.H + h) * p.Lq + q;
                lse2 = q_ok ? __ldg(p.lse + st) * AF_LOG2E : INFINITY;  // +inf -> P = 0 for padded query rows
                dlt = acc * p.scale;                                     // pre-scaled: dS = P·(scale·dP − scale·delta)
                my_lse[i * AF_T + row] = lse2;
                my_dlt[i * AF_T + row] = dlt;
                if (half == 0 && q_ok && p.delta) p.delta[st] = acc;
            } else {
                lse2 = my_lse[i * AF_T + row];
                dlt = my_dlt[i * AF_T + row];
            }
            // P and dS are formed IN PLACE in the registers that hold S and dP, before anything is waited for: the
            // exponentials of tile t overlap the gradient MMAs of tile t-1
            const int kvalid = p.Lk - j * AF_T;  // >= 1
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
#pragma unroll
                for (int idx = 0; idx < 32; ++idx) {
                    const float x = fast_exp2(fmaf(__uint_as_float(vs[cc][idx]), sl2, -lse2));
                    const float g = fmaf(__uint_as_float(vp[cc][idx]), p.scale, -dlt);
                    vs[cc][idx] = __float_as_uint(x);
                    vp[cc][idx] = __float_as_uint(x * g);
                }
            }
            if (kvalid < AF_T) {  // ragged last key tile (warp-uniform): keys beyond Lk contribute nothing
#pragma unroll
                for (int cc = 0; cc < 2; ++cc)
#pragma unroll
                    for (int idx = 0; idx < 32; ++idx)
                        if ((2 * half + cc) * 32 + idx >= kvalid) {
                            vs[cc][idx] = 0u;
                            vp[cc][idx] = 0u;
                        }
            }
            if (t > 0) {
                mbar_wait(g_done, (t - 1) & 1);  // gradient MMAs of tile t-1 retired: P / dS tiles may be overwritten
                tc_fence_after();
            }
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = 2 * half + cc;  // 32-column chunk of the 128-wide tile
                uint8_t* slabP = sP + (c >> 1) * (AF_T * 128);
                uint8_t* slabD = sDS + (c >> 1) * (AF_T * 128);
#pragma unroll
                for (int qd = 0; qd < 4; ++qd) {
                    uint4 up, ud;
                    up.x = pack_bf16x2(__uint_as_float(vs[cc][qd * 8 + 0]), __uint_as_float(vs[cc][qd * 8 + 1]));
                    up.y = pack_bf16x2(__uint_as_float(vs[cc][qd * 8 + 2]), __uint_as_float(vs[cc][qd * 8 + 3]));
                    up.z = pack_bf16x2(__uint_as_float(vs[cc][qd * 8 + 4]), __uint_as_float(vs[cc][qd * 8 + 5]));
                    up.w = pack_bf16x2(__uint_as_float(vs[cc][qd * 8 + 6]), __uint_as_float(vs[cc][qd * 8 + 7]));
                    ud.x = pack_bf16x2(__uint_as_float(vp[cc][qd * 8 + 0]), __uint_as_float(vp[cc][qd * 8 + 1]));
                    ud.y = pack_bf16x2(__uint_as_float(vp[cc][qd * 8 + 2]), __uint_as_float(vp[cc][qd * 8 + 3]));
                    ud.z = pack_bf16x2(__uint_as_float(vp[cc][qd * 8 + 4]), __uint_as_float(vp[cc][qd * 8 + 5]));
                    ud.w = pack_bf16x2(__uint_as_float(vp[cc][qd * 8 + 6]), __uint_as_float(vp[cc][qd * 8 + 7]));
                    const uint32_t off = swz_off<128>(row, (c & 1) * 4 + qd);
                    *reinterpret_cast<uint4*>(slabP + off) = up;
                    *reinterpret_cast<uint4*>(slabD + off) = ud;
                }
            }
            // dV / dK of the previous key tile are final (g_done above); tile t's gradient MMAs, which restart them,
            // are only issued after this thread's p_full arrival below
            if (t > 0 && i == 0) read_out_dkv(j - 1);
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(p_full);
        }
        mbar_wait(g_done, (ntiles - 1) & 1);
        tc_fence_after();
        read_out_dkv(nkt - 1);
        // dQ of every query tile: (tile, part) pairs alternate between the two warps of a lane quarter
        uint32_t v[32];
        for (int idx = half; idx < nqt * NP; idx += 2) {
            const int i = idx / NP, part = idx - i * NP;
            const int q = i * AF_T + row;
            tmem_ld_32x32b_x32(tmem_dQ + lane_off + idx * AF_D, v);
            tmem_ld_wait();
            if (q < p.Lq) {
                const long long ldg = part ? C : p.ld_dq;
                __nv_bfloat16* dst = (part ? p.dq2 : p.dq1) + ((long long)b * p.Lq + q) * ldg + h * AF_D;
                af_store_row32(dst, v);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == AF_MMA) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

static int make_seq_tmap_f(CUtensorMap* tm, const void* base, int rowlen, int L, int B, long long ld, long long bs) {
    const bool bcast = (bs == 0) || (B == 1);
    uint64_t dims[3] = {(uint64_t)rowlen, (uint64_t)L, (uint64_t)(bcast ? 1 : B)};
    uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)(bcast ? (uint64_t)ld * 2 * L : (uint64_t)bs * 2)};
    uint32_t box[3] = {AF_D, AF_T, 1};
    return make_tmap_bf16(tm, base, 3, dims, str, box, 64);
}

bool attn_bwd_fused_supported(const fod_attn_bwd_args* g) {
    const fod_attn_args* a = &g->fwd;
    const int nqt = (a->Lq + AF_T - 1) / AF_T;
    return nqt <= (a->q2 ? FusedBwdSmem<2>::kMaxQT : FusedBwdSmem<1>::kMaxQT);
}

// Arguments were validated by fod_attn_bwd.
int launch_attn_bwd_fused(const fod_attn_bwd_args* g, cudaStream_t st) {
    const fod_attn_args* a = &g->fwd;
    const int two = a->q2 != nullptr;
    const int rowlen = a->H * AF_D;
    CUtensorMap tq1, tk1, tq2, tk2, tv, tdo;
    if (make_seq_tmap_f(&tq1, a->q1, rowlen, a->Lq, a->B, a->ldq1, a->bsq1)) return fod_fail(FOD_E_TMAP, "attn_bwd: q1");
    if (make_seq_tmap_f(&tk1, a->k1, rowlen, a->Lk, a->B, a->ldk1, a->bsk1)) return fod_fail(FOD_E_TMAP, "attn_bwd: k1");
    if (make_seq_tmap_f(&tv, a->v, rowlen, a->Lk, a->B, a->ldv, a->bsv)) return fod_fail(FOD_E_TMAP, "attn_bwd: v");
    if (make_seq_tmap_f(&tdo, g->dout, rowlen, a->Lq, a->B, g->lddo, a->B > 1 ? g->bsdo : 0))
        return fod_fail(FOD_E_TMAP, "attn_bwd: dout");
    if (two) {
        if (make_seq_tmap_f(&tq2, a->q2, rowlen, a->Lq, a->B, a->ldq2, a->bsq2)) return fod_fail(FOD_E_TMAP, "attn_bwd: q2");
        if (make_seq_tmap_f(&tk2, a->k2, rowlen, a->Lk, a->B, a->ldk2, a->bsk2)) return fod_fail(FOD_E_TMAP, "attn_bwd: k2");
    } else {
        tq2 = tq1;
        tk2 = tk1;
    }
    AttnBwdFusedParams p;
    p.lse = a->lse;
    p.delta = g->delta;
    p.out = reinterpret_cast<const __nv_bfloat16*>(a->out);
    p.ldo = a->ldo;
    p.bso = a->bso;
    p.dq1 = reinterpret_cast<__nv_bfloat16*>(g->dq1);
    p.dq2 = reinterpret_cast<__nv_bfloat16*>(g->dq2);
    p.dk1 = reinterpret_cast<__nv_bfloat16*>(g->dk1);
    p.dk2 = reinterpret_cast<__nv_bfloat16*>(g->dk2);
    p.dv = reinterpret_cast<__nv_bfloat16*>(g->dv);
    p.ld_dq = g->ld_dq ? g->ld_dq : rowlen;
    p.ld_dk = g->ld_dk ? g->ld_dk : rowlen;
    p.B = a->B; p.H = a->H; p.Lq = a->Lq; p.Lk = a->Lk;
    auto has_b = [&](long long bs) { return (bs != 0 && a->B > 1) ? 1 : 0; };
    p.bm_q1 = has_b(a->bsq1); p.bm_k1 = has_b(a->bsk1);
    p.bm_q2 = two ? has_b(a->bsq2) : 0; p.bm_k2 = two ? has_b(a->bsk2) : 0;
    p.bm_v = has_b(a->bsv);
    p.scale = a->scale;
    static bool attr[2] = {false, false};
    if (!attr[two]) {
        if (two) cudaFuncSetAttribute(attn_bwd_fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, FusedBwdSmem<2>::kTotal);
        else cudaFuncSetAttribute(attn_bwd_fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, FusedBwdSmem<1>::kTotal);
        attr[two] = true;
    }
    dim3 grid(a->H, a->B);
    if (two) fod_launch(attn_bwd_fused_kernel<2>, grid, dim3(AF_THREADS), FusedBwdSmem<2>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
    else fod_launch(attn_bwd_fused_kernel<1>, grid, dim3(AF_THREADS), FusedBwdSmem<1>::kTotal, st, tq1, tk1, tq2, tk2, tv, tdo, p);
    return fod_check_launch("attn_bwd_fused_kernel");
}

}  // namespace fod
