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
//   delta = rowsum(dO ⊙ O) is computed by the softmax warps in a pre-pass that overlaps the first loads (no separate
//   kernel).
//
// TMEM: S 128 + dP 128 + dV 32 + dK 32·NP + dQ 32·NP·nqt  <= 512 columns  =>  nqt <= 6 (one part), nqt <= 2 (two parts).
// Longer sequences take the two-kernel path of attn_bwd_sm100.cu.
// Bound: MUFU ex2 (one exponential per 160 tensor flop at head dim 32): ~1024 clk per 128x128 tile pair and SM.
// Replaces the autograd of ConditionalDETR/models/attention.py:350-372 and of torch.nn.MultiheadAttention's core
// (future_od/models/transformer.py:404,417).
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"
#include "dropout.cuh"

#ifndef FOD_AF_TRACE
#define FOD_AF_TRACE 0
#endif

namespace fod {

#if FOD_AF_TRACE
// Timeline probe (debug builds only): clock64 stamps of CTA 0 — [role][tile][event]
__device__ long long g_af_trace[3 * 16 * 8];
#define AF_STAMP(role, tile, ev)                                                                   \
    do {                                                                                           \
        if (blockIdx.x == 0 && blockIdx.y == 0 && (tile) < 16) g_af_trace[((role) * 16 + (tile)) * 8 + (ev)] = clock64(); \
    } while (0)
#else
#define AF_STAMP(role, tile, ev) do {} while (0)
#endif

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
    static constexpr int kQS = 2;                           // Q / dO stages
    static constexpr int kQ = kV + 2 * AF_TILE;             // Q parts, kQS stages
    static constexpr int kDO = kQ + kQS * NP * AF_TILE;     // dO, kQS stages
    // one-part attention double-buffers the P / dS tiles: the softmax warps of tile t only wait for the gradient MMAs
    // of tile t-2 (those MMAs are operand-fetch bound, ~65 clk each whatever their width: 24 of them outlast the
    // exponentials of the next tile's first half); with two parts the second buffer does not fit
    static constexpr int kPBuf = NP == 1 ? 2 : 1;
    static constexpr int kP = kDO + kQS * AF_TILE;          // P tiles
    static constexpr int kDS = kP + kPBuf * AF_PT;          // dS tiles
    static constexpr int kOut = kDS + kPBuf * AF_PT;        // 3 staging tiles [128][32] bf16 of the TMA gradient stores
    static constexpr int kStat = kOut + 3 * AF_TILE;        // [lse2 | scale*delta][kMaxQT*128] fp32
    static constexpr int kBars = kStat + 2 * kMaxQT * AF_T * 4;
    static constexpr int kTotal = kBars + 256 + 1024;       // + alignment slack
};

struct AttnBwdFusedParams {
    const float* lse;            // (B,H,Lq)
    float* delta;                // (B,H,Lq) optional output (kept for API parity with the two-kernel path)
    const __nv_bfloat16* out;    // O (B,Lq,H*32)
    const __nv_bfloat16* dout;   // dO (B,Lq,H*32)
    long long ldo, bso, lddo, bsdo;
    __nv_bfloat16 *dq1, *dq2, *dk1, *dk2, *dv;
    long long ld_dq, ld_dk;
    int B, H, Lq, Lk;
    int bm_q1, bm_k1, bm_q2, bm_k2, bm_v;
    float scale;
    DropoutCfg drop;             // DROP kernels: the forward call's dropout on P, regenerated here (DROP == 1)
    const uint8_t* keep_bits;    // ... or read from the bits the forward call used (DROP == 2, fod_attn_keep_bits)
    long long keep_pitch;
    // Query groups (grid.z > 1, one-part attention longer than kMaxQT query tiles): CTA z owns query tiles
    // [z*kMaxQT, z*kMaxQT + kMaxQT) and ALL key tiles.  dQ is complete per CTA; dK / dV are partial sums over the
    // group's queries and are reduced with fp32 vector atomics into these (B,H,Lk,32) workspaces (zeroed by the host
    // side, converted to bf16 by attn_ws_to_bf16_kernel afterwards).  nullptr: direct bf16 TMA stores.
    float* ws_dk;
    float* ws_dv;
};

// One [128][32] fp32 accumulator tile, TMEM -> bf16 -> swizzled staging tile -> TMA store.  Executed by the four warps
// of a `half` (128 threads, named barrier 2 + half); `issuer` is the one thread of the half that owns its bulk groups.
__device__ __forceinline__ void af_store_tile(uint32_t taddr_lane, uint8_t* slot, const CUtensorMap* tm, int row0, int row,
                                              int half, bool issuer, int h, int b) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(taddr_lane, v);
    if (issuer) bulk_wait_read0();  // the previous store out of a slot of this half has finished reading it
    asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint4 u;
        u.x = pack_bf16x2(__uint_as_float(v[c * 8 + 0]), __uint_as_float(v[c * 8 + 1]));
        u.y = pack_bf16x2(__uint_as_float(v[c * 8 + 2]), __uint_as_float(v[c * 8 + 3]));
        u.z = pack_bf16x2(__uint_as_float(v[c * 8 + 4]), __uint_as_float(v[c * 8 + 5]));
        u.w = pack_bf16x2(__uint_as_float(v[c * 8 + 6]), __uint_as_float(v[c * 8 + 7]));
        *reinterpret_cast<uint4*>(slot + swz_off<64>(row, c)) = u;
    }
    fence_proxy_async_smem();
    asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
    if (issuer) {
        tma_store_3d(tm, slot, h * AF_D, row0, b);
        bulk_commit();
    }
}

// One [128][32] fp32 accumulator tile added into a (.., Lk, 32) fp32 workspace: a thread owns one key row (128 contiguous
// bytes) and issues eight 16-byte vector reductions.
__device__ __forceinline__ void af_red_tile(uint32_t taddr_lane, float* ws_row, bool valid) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(taddr_lane, v);
    tmem_ld_wait();
    if (valid) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
            atomicAdd(reinterpret_cast<float4*>(ws_row) + c,
                      make_float4(__uint_as_float(v[c * 4]), __uint_as_float(v[c * 4 + 1]), __uint_as_float(v[c * 4 + 2]),
                                  __uint_as_float(v[c * 4 + 3])));
    }
}

// DROP: out = dropout(P)·V in the forward pass (mask m, scale c = 1/(1-p)): dV takes P·m·c, dP = (dO·Vᵀ)·m·c, and
// dS = P ⊙ (dP − delta) with the UNdropped P; delta = rowsum(dO ⊙ O) is unchanged.
template <int NP, int DROP>
__global__ void __launch_bounds__(AF_THREADS, 1)
attn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tmQ1, const __grid_constant__ CUtensorMap tmK1,
                      const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmK2,
                      const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                      const __grid_constant__ CUtensorMap tmDQ1, const __grid_constant__ CUtensorMap tmDQ2,
                      const __grid_constant__ CUtensorMap tmDK1, const __grid_constant__ CUtensorMap tmDK2,
                      const __grid_constant__ CUtensorMap tmDV, const AttnBwdFusedParams p) {
    using L = FusedBwdSmem<NP>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* kv_full = bars + 0;    // [2]
    uint64_t* kv_empty = bars + 2;   // [2]
    uint64_t* qdo_full = bars + 4;   // [kQS <= 3]
    uint64_t* qdo_empty = bars + 7;  // [kQS <= 3]
    uint64_t* s_full = bars + 10;    // S / dP of tile t are complete in TMEM
    uint64_t* s_free = bars + 11;    // ... and have been copied into registers by all 256 softmax threads
    uint64_t* p_full = bars + 12;    // P / dS of tile t are in shared memory
    // the gradient MMAs of tile t have retired (P / dS / operand stages reusable): alternates between two barriers, as
    // a parity wait can only tell the latest phase from the one before and the softmax warps look two tiles back
    uint64_t* g_done = bars + 13;    // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);
    float* stat = reinterpret_cast<float*>(smem + L::kStat);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    if (threadIdx.x == 0) AF_STAMP(2, 0, 0);  // kernel entry
    const int qt0 = blockIdx.z * L::kMaxQT;                                   // first query tile of this CTA's group
    const int nqt = min(L::kMaxQT, (p.Lq + AF_T - 1) / AF_T - qt0);           // its query tiles (local index i)
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
        }
        for (int s = 0; s < L::kQS; ++s) {
            mbar_init(qdo_full + s, 1);
            mbar_init(qdo_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(s_free, 256);
        mbar_init(p_full, 256);
        mbar_init(g_done, 1);
        mbar_init(g_done + 1, 1);
        fence_barrier_init();
    }
    if (warp == AF_MMA) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) AF_STAMP(2, 0, 1);  // barriers initialised, TMEM allocated
    pdl_launch_dependents();
    pdl_wait();
    if (threadIdx.x == 0) AF_STAMP(2, 0, 2);  // predecessor grid complete
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
                    const int s = t % L::kQS;
                    mbar_wait(qdo_empty + s, ((t / L::kQS) & 1) ^ 1);
                    mbar_arrive_expect_tx(qdo_full + s, (NP + 1) * AF_TILE);
                    uint8_t* sq = smem + L::kQ + s * NP * AF_TILE;
                    tma_load_3d(sq, &tmQ1, qdo_full + s, h * AF_D, (qt0 + i) * AF_T, b * p.bm_q1);
                    if (NP == 2) tma_load_3d(sq + AF_TILE, &tmQ2, qdo_full + s, h * AF_D, (qt0 + i) * AF_T, b * p.bm_q2);
                    tma_load_3d(smem + L::kDO + s * AF_TILE, &tmDO, qdo_full + s, h * AF_D, (qt0 + i) * AF_T, b);
                }
            }
        }
    } else if (warp == AF_MMA) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(AF_T, AF_T, 0, 0);   // S, dP: both operands K-major
            constexpr uint32_t idesc_q = umma_idesc_bf16(AF_T, AF_D, 0, 1);   // dQ: A = dS K-major, B = K_j MN-major
            constexpr uint32_t idesc_k = umma_idesc_bf16(AF_T, AF_D, 1, 1);   // dV, dK: A = Pᵀ / dSᵀ (MN-major view)
            // Descriptors of stage 0 of every operand buffer, built once; per tile / per instruction they are stepped
            // with umma_desc_adv (one add each).
            //   K-major 64-byte rows (SWIZZLE_64B): LBO 16, SBO 512, 16 contraction elements = +32 B
            //   MN-major B from the same tiles ([row = contraction][32 d]): LBO 4096, SBO 512, 16 rows = +1024 B
            //   P / dS tile as K-major A (rows = queries, SWIZZLE_128B slabs of 64 keys): LBO 16, SBO 1024, +32 B / +16 KB
            //   P / dS tile as MN-major A (contraction = query rows): LBO 16 KB (next 64 keys), SBO 1024, 16 rows = +2048 B
            const uint64_t dK_k = umma_smem_desc(smem_u32(smem + L::kK), 16, 512, 4);
            const uint64_t dV_k = umma_smem_desc(smem_u32(smem + L::kV), 16, 512, 4);
            const uint64_t dQ_k = umma_smem_desc(smem_u32(smem + L::kQ), 16, 512, 4);
            const uint64_t dDO_k = umma_smem_desc(smem_u32(smem + L::kDO), 16, 512, 4);
            const uint64_t dK_mn = umma_smem_desc(smem_u32(smem + L::kK), 4096, 512, 4);
            const uint64_t dQ_mn = umma_smem_desc(smem_u32(smem + L::kQ), 4096, 512, 4);
            const uint64_t dDO_mn = umma_smem_desc(smem_u32(smem + L::kDO), 4096, 512, 4);
            const uint64_t dP_mn = umma_smem_desc(smem_u32(smem + L::kP), AF_T * 128, 1024, 2);
            const uint64_t dDS_mn = umma_smem_desc(smem_u32(smem + L::kDS), AF_T * 128, 1024, 2);
            const uint64_t dDS_k = umma_smem_desc(smem_u32(smem + L::kDS), 16, 1024, 2);
            auto issue_SP = [&](int t) {
                const int j = t / nqt, i = t - j * nqt;
                const int sj = j & 1, s = t % L::kQS;
                if (i == 0) mbar_wait(kv_full + sj, (j >> 1) & 1);
                mbar_wait(qdo_full + s, (t / L::kQS) & 1);
                tc_fence_after();
                AF_STAMP(1, t, 1);
                const uint64_t k_d = umma_desc_adv(dK_k, sj * NP * AF_TILE), v_d = umma_desc_adv(dV_k, sj * AF_TILE);
                const uint64_t q_d = umma_desc_adv(dQ_k, s * NP * AF_TILE), do_d = umma_desc_adv(dDO_k, s * AF_TILE);
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int k = 0; k < AF_D / 16; ++k)
                        umma_bf16_ss(tmem_S, umma_desc_adv(q_d, part * AF_TILE + k * 32),
                                     umma_desc_adv(k_d, part * AF_TILE + k * 32), idesc_s, (part | k) ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < AF_D / 16; ++k)
                    umma_bf16_ss(tmem_dP, umma_desc_adv(do_d, k * 32), umma_desc_adv(v_d, k * 32), idesc_s, k ? 1u : 0u);
                umma_commit(s_full);
            };
            issue_SP(0);
            for (int t = 0; t < ntiles; ++t) {
                const int j = t / nqt, i = t - j * nqt;
                const int sj = j & 1, s = t % L::kQS;
                if (t + 1 < ntiles) {
                    // S(t) / dP(t) live in the softmax warps' registers: the next pair is computed while they work
                    mbar_wait(s_free, t & 1);
                    tc_fence_after();
                    AF_STAMP(1, t + 1, 0);
                    issue_SP(t + 1);
                    AF_STAMP(1, t + 1, 2);
                }
                mbar_wait(p_full, t & 1);
                tc_fence_after();
                AF_STAMP(1, t, 3);
                const uint64_t k_mn = umma_desc_adv(dK_mn, sj * NP * AF_TILE);
                const uint64_t q_mn = umma_desc_adv(dQ_mn, s * NP * AF_TILE), do_mn = umma_desc_adv(dDO_mn, s * AF_TILE);
                // (every MMA below is issued unconditionally: individually predicated tcgen05.mma in these unrolled loops
                // — an attempt to stop the contraction at a ragged tile's last valid row — produced wrong dQ and faults)
                const uint32_t acc_i = i ? 1u : 0u, acc_j = j ? 1u : 0u;
                const uint32_t pb = (t % L::kPBuf) * AF_PT;  // which P / dS buffer this tile used
                const uint64_t p_mn = umma_desc_adv(dP_mn, pb), ds_mn = umma_desc_adv(dDS_mn, pb), ds_k = umma_desc_adv(dDS_k, pb);
                // dV_j (+)= Pᵀ·dO_i: contraction over the 128 queries, 16 per instruction
#pragma unroll
                for (int kk = 0; kk < AF_T / 16; ++kk)
                    umma_bf16_ss(tmem_dV, umma_desc_adv(p_mn, kk * 2048), umma_desc_adv(do_mn, kk * 1024), idesc_k,
                                     kk ? 1u : acc_i);
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int kk = 0; kk < AF_T / 16; ++kk)
                        umma_bf16_ss(tmem_dK + part * AF_D, umma_desc_adv(ds_mn, kk * 2048),
                                         umma_desc_adv(q_mn, part * AF_TILE + kk * 1024), idesc_k, kk ? 1u : acc_i);
                // dQ_i (+)= dS·K_j: contraction over the 128 keys
                const uint32_t tq = tmem_dQ + i * NP * AF_D;
#pragma unroll
                for (int part = 0; part < NP; ++part)
#pragma unroll
                    for (int kk = 0; kk < AF_T / 16; ++kk)
                        umma_bf16_ss(tq + part * AF_D, umma_desc_adv(ds_k, (kk >> 2) * (AF_T * 128) + (kk & 3) * 32),
                                         umma_desc_adv(k_mn, part * AF_TILE + kk * 1024), idesc_q, kk ? 1u : acc_j);
                umma_commit(g_done + (t & 1));
                AF_STAMP(1, t, 4);
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
        float* s_lse = stat;                        // lse·log2(e) per query row (+inf for padded rows: P = 0)
        float* s_dlt = stat + L::kMaxQT * AF_T;     // scale·delta per query row
        uint8_t* sP = smem + L::kP;
        uint8_t* sDS = smem + L::kDS;

        // Pre-pass (overlaps the first TMA loads and score MMAs): delta = rowsum(dO ⊙ O) and lse of EVERY query row of
        // this (batch, head).  Four lanes share a row (16 bytes each: a warp-wide load touches 8 lines instead of 32),
        // all loads of all query tiles are issued before the first use, partial dot products meet by shuffle.
        {
            const __nv_bfloat16* o_base = p.out + (long long)b * p.bso + h * AF_D;
            const __nv_bfloat16* do_base = p.dout + (long long)b * p.bsdo + h * AF_D;
            const int tid = threadIdx.x;          // 0..255
            const int r0 = tid >> 2, ch = tid & 3;
            uint4 a4[L::kMaxQT][2], d4[L::kMaxQT][2];
            float l2[L::kMaxQT][2];
            const long long st0 = ((long long)b * p.H + h) * p.Lq;
#pragma unroll
            for (int i = 0; i < L::kMaxQT; ++i)
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int q = (qt0 + i) * AF_T + u * 64 + r0;
                    const bool ok = i < nqt && q < p.Lq;
                    a4[i][u] = ok ? __ldg(reinterpret_cast<const uint4*>(o_base + (long long)q * p.ldo) + ch) : make_uint4(0, 0, 0, 0);
                    d4[i][u] = ok ? __ldg(reinterpret_cast<const uint4*>(do_base + (long long)q * p.lddo) + ch) : make_uint4(0, 0, 0, 0);
                    l2[i][u] = (ok && ch == 0) ? __ldg(p.lse + st0 + q) : INFINITY;  // +inf: P = 0 for padded rows
                }
#pragma unroll
            for (int i = 0; i < L::kMaxQT; ++i)
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (i >= nqt) break;  // warp-uniform
                    const int rr = u * 64 + r0, q = (qt0 + i) * AF_T + rr;
                    const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a4[i][u]);
                    const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&d4[i][u]);
                    float acc = 0.f;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 x = __bfloat1622float2(ha[e]), y = __bfloat1622float2(hd[e]);
                        acc = fmaf(x.x, y.x, acc);
                        acc = fmaf(x.y, y.y, acc);
                    }
                    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                    if (ch == 0) {
                        s_lse[i * AF_T + rr] = l2[i][u] * AF_LOG2E;
                        s_dlt[i * AF_T + rr] = acc * p.scale;  // pre-scaled: dS = P·(scale·dP − scale·delta)
                        if (q < p.Lq && p.delta) p.delta[st0 + q] = acc;
                    }
                }
            asm volatile("bar.sync 1, 256;" ::: "memory");  // the 8 softmax warps
        }
        if (threadIdx.x == 0) AF_STAMP(2, 0, 3);  // delta pre-pass done

        // Gradient tiles leave through TMA: the four warps of a `half` (one per TMEM lane quarter) write their rows of
        // the [128][32] accumulator into a swizzled staging tile and one thread issues the bulk store (rows beyond the
        // sequence end are clipped by the tensor map).  Row-per-thread global stores would touch 32 lines per warp
        // instruction and cost ~1.5 us per CTA in the load/store unit.
        uint8_t* my_slot = smem + L::kOut + half * AF_TILE;     // half 0: slots 0 and 2, half 1: slot 1
        const bool issuer = quarter == 0 && lane == 0;
        // (a force-inlined function, not a lambda: out of line, the tensor-map argument would be a run-time pointer and
        // the by-reference captures would live in local memory)
#define AF_STORE_TILE(taddr, slot, tm, row0) \
    af_store_tile((taddr) + lane_off, (slot), (tm), (row0), row, half, issuer, h, b)
        auto read_out_dkv = [&](int jj) {  // dV / dK of key tile jj
            if (NP == 1 && p.ws_dk) {  // query groups: partial sums into the fp32 workspaces
                const int key = jj * AF_T + row;
                float* dst = (half == 0 ? p.ws_dv : p.ws_dk) + (((long long)b * p.H + h) * p.Lk + key) * AF_D;
                af_red_tile((half == 0 ? tmem_dV : tmem_dK) + lane_off, dst, key < p.Lk);
                return;
            }
            if (half == 0) {
                AF_STORE_TILE(tmem_dV, my_slot, &tmDV, jj * AF_T);
                if (NP == 2) AF_STORE_TILE(tmem_dK + AF_D, smem + L::kOut + 2 * AF_TILE, &tmDK2, jj * AF_T);
            } else {
                AF_STORE_TILE(tmem_dK, my_slot, &tmDK1, jj * AF_T);
            }
        };

        uint32_t vs[32], vp[32];
        PhiloxKey key = {0u, 0u, 0u, 0u};
        if (DROP == 1) key = dropout_key(p.drop);
        const unsigned long long lkp8 = (unsigned long long)((p.Lk + 7) >> 3);  // Philox groups per (padded) score row
        for (int t = 0; t < ntiles; ++t) {
            const int j = t / nqt, i = t - j * nqt;
            const float lse2 = s_lse[i * AF_T + row];
            const float dlt = s_dlt[i * AF_T + row];
            const unsigned long long g_row =
                (((unsigned long long)b * p.H + h) * p.Lq + (unsigned long long)min((qt0 + i) * AF_T + row, p.Lq - 1)) * lkp8;
            const int kvalid = p.Lk - j * AF_T;  // >= 1
            uint2 kb2 = make_uint2(0u, 0u);      // DROP == 2: the 64 keep bits of this thread's two 32-column chunks
            if (DROP == 2)
                kb2 = __ldg(reinterpret_cast<const uint2*>(
                                p.keep_bits + (((long long)b * p.H + h) * p.Lq + min((qt0 + i) * AF_T + row, p.Lq - 1)) * p.keep_pitch) +
                            j * 2 + half);
            if (threadIdx.x == 0) AF_STAMP(0, t, 0);
            mbar_wait(s_full, t & 1);
            tc_fence_after();
            if (threadIdx.x == 0) AF_STAMP(0, t, 1);
            // Two passes of 32 columns (register budget: 168 per thread at 320 threads): P and dS are formed, packed to
            // bf16 and stored per pass.  The exponentials of the first pass run before anything else is waited for, so
            // they overlap the gradient MMAs of tile t-1; S / dP in TMEM are released after the second load, so the
            // next score pair is computed while this one is still being exponentiated.
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = 2 * half + cc;  // 32-column chunk of the 128-wide tile
                tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, vs);
                tmem_ld_32x32b_x32(tmem_dP + lane_off + c * 32, vp);
                tmem_ld_wait();
                if (cc == 1) {
                    tc_fence_before();
                    mbar_arrive(s_free);
                }
                if (threadIdx.x == 0) AF_STAMP(0, t, cc == 0 ? 2 : 5);
                if (DROP) {
#pragma unroll
                    for (int qd = 0; qd < 4; ++qd) {
                        Keep8 keep = {{0u, 0u, 0u, 0u}, 0u};
                        if (DROP == 1) keep = dropout_draw8(key, g_row + (unsigned long long)(j * (AF_T / 8) + c * 4 + qd), p.drop.thresh);
                        const uint32_t kw = (cc == 0 ? kb2.x : kb2.y) >> (8 * qd);
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const int idx = qd * 8 + e;
                            const bool kept = DROP == 1 ? keep[e] : ((kw >> e) & 1u) != 0u;
                            const float mk = kept ? p.drop.scale : 0.f;
                            const float x = fast_exp2(fmaf(__uint_as_float(vs[idx]), sl2, -lse2));
                            const float g = fmaf(__uint_as_float(vp[idx]) * mk, p.scale, -dlt);
                            vs[idx] = __float_as_uint(x * mk);  // dropout(P): the operand of dV
                            vp[idx] = __float_as_uint(x * g);   // dS from the undropped P
                        }
                    }
                } else {
#pragma unroll
                    for (int idx = 0; idx < 32; ++idx) {
                        const float x = fast_exp2(fmaf(__uint_as_float(vs[idx]), sl2, -lse2));
                        const float g = fmaf(__uint_as_float(vp[idx]), p.scale, -dlt);
                        vs[idx] = __float_as_uint(x);
                        vp[idx] = __float_as_uint(x * g);
                    }
                }
                if (kvalid < AF_T) {  // ragged last key tile (warp-uniform): keys beyond Lk contribute nothing
#pragma unroll
                    for (int idx = 0; idx < 32; ++idx)
                        if (c * 32 + idx >= kvalid) {
                            vs[idx] = 0u;
                            vp[idx] = 0u;
                        }
                }
                if (cc == 0) {
                    if (threadIdx.x == 0) AF_STAMP(0, t, 3);
                    if (t >= L::kPBuf) {  // the gradient MMAs that read this P / dS buffer last have retired
                        const int tp = t - L::kPBuf;
                        mbar_wait(g_done + (tp & 1), (tp >> 1) & 1);
                        tc_fence_after();
                    }
                    if (threadIdx.x == 0) AF_STAMP(0, t, 4);
                }
                uint8_t* slabP = sP + (t % L::kPBuf) * AF_PT + (c >> 1) * (AF_T * 128);
                uint8_t* slabD = sDS + (t % L::kPBuf) * AF_PT + (c >> 1) * (AF_T * 128);
#pragma unroll
                for (int qd = 0; qd < 4; ++qd) {
                    uint4 up, ud;
                    up.x = pack_bf16x2(__uint_as_float(vs[qd * 8 + 0]), __uint_as_float(vs[qd * 8 + 1]));
                    up.y = pack_bf16x2(__uint_as_float(vs[qd * 8 + 2]), __uint_as_float(vs[qd * 8 + 3]));
                    up.z = pack_bf16x2(__uint_as_float(vs[qd * 8 + 4]), __uint_as_float(vs[qd * 8 + 5]));
                    up.w = pack_bf16x2(__uint_as_float(vs[qd * 8 + 6]), __uint_as_float(vs[qd * 8 + 7]));
                    ud.x = pack_bf16x2(__uint_as_float(vp[qd * 8 + 0]), __uint_as_float(vp[qd * 8 + 1]));
                    ud.y = pack_bf16x2(__uint_as_float(vp[qd * 8 + 2]), __uint_as_float(vp[qd * 8 + 3]));
                    ud.z = pack_bf16x2(__uint_as_float(vp[qd * 8 + 4]), __uint_as_float(vp[qd * 8 + 5]));
                    ud.w = pack_bf16x2(__uint_as_float(vp[qd * 8 + 6]), __uint_as_float(vp[qd * 8 + 7]));
                    const uint32_t off = swz_off<128>(row, (c & 1) * 4 + qd);
                    *reinterpret_cast<uint4*>(slabP + off) = up;
                    *reinterpret_cast<uint4*>(slabD + off) = ud;
                }
            }
            // dV / dK of the previous key tile are final once tile t-1's gradient MMAs have retired; tile t's, which
            // restart them, are only issued after this thread's p_full arrival below
            if (t > 0 && i == 0) {
                mbar_wait(g_done + ((t - 1) & 1), ((t - 1) >> 1) & 1);
                tc_fence_after();
                read_out_dkv(j - 1);
            }
            if (threadIdx.x == 0) AF_STAMP(0, t, 6);
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(p_full);
            if (threadIdx.x == 0) AF_STAMP(0, t, 7);
        }
        if (threadIdx.x == 0) AF_STAMP(2, 0, 4);  // main loop done
        mbar_wait(g_done + ((ntiles - 1) & 1), ((ntiles - 1) >> 1) & 1);
        tc_fence_after();
        if (threadIdx.x == 0) AF_STAMP(2, 0, 5);  // last gradient MMAs retired
        read_out_dkv(nkt - 1);
        // dQ of every query tile: (tile, part) pairs alternate between the two halves
        for (int idx = half; idx < nqt * NP; idx += 2) {
            const int i = idx / NP, part = idx - i * NP;
            // (a run-time choice between two tensor-map POINTERS would make the compiler copy the __grid_constant__
            // maps to local memory, which the TMA unit cannot read: branch instead)
            if (NP == 2 && part) AF_STORE_TILE(tmem_dQ + idx * AF_D, my_slot, &tmDQ2, (qt0 + i) * AF_T);
            else AF_STORE_TILE(tmem_dQ + idx * AF_D, my_slot, &tmDQ1, (qt0 + i) * AF_T);
        }
        if (issuer) bulk_wait0();  // the staging tiles live in this CTA's shared memory: drain before it is released
        tc_fence_before();
        if (threadIdx.x == 0) AF_STAMP(2, 0, 6);  // gradients stored
    }
    __syncthreads();
    if (warp == AF_MMA) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
    if (threadIdx.x == 0) AF_STAMP(2, 0, 7);
}

// Query-group launches: zero the fp32 dK / dV workspaces, and afterwards write them as bf16 into the (strided) outputs.
__global__ void attn_ws_zero_kernel(float4* __restrict__ ws, long long n4) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
        ws[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}
// one thread per 8 consecutive channels of one (batch, key, head)
__global__ void attn_ws_to_bf16_kernel(const float* __restrict__ ws_dk, const float* __restrict__ ws_dv,
                                       __nv_bfloat16* __restrict__ dk, __nv_bfloat16* __restrict__ dv, long long ld, int B,
                                       int H, int Lk) {
    const long long n = (long long)B * Lk * H * 4;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(t & 3);
        const int h = (int)((t >> 2) % H);
        const long long bk = (t >> 2) / H;  // b * Lk + key
        const int key = (int)(bk % Lk);
        const int b = (int)(bk / Lk);
        const long long src = ((((long long)b * H + h) * Lk + key) * AF_D) + c * 8;
        const long long dst = bk * ld + h * AF_D + c * 8;
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            const float4 x = *reinterpret_cast<const float4*>((w ? ws_dv : ws_dk) + src);
            const float4 y = *reinterpret_cast<const float4*>((w ? ws_dv : ws_dk) + src + 4);
            uint4 u;
            u.x = pack_bf16x2(x.x, x.y);
            u.y = pack_bf16x2(x.z, x.w);
            u.z = pack_bf16x2(y.x, y.y);
            u.w = pack_bf16x2(y.z, y.w);
            *reinterpret_cast<uint4*>((w ? dv : dk) + dst) = u;
        }
    }
}

static int make_seq_tmap_f(CUtensorMap* tm, const void* base, int rowlen, int L, int B, long long ld, long long bs) {
    const bool bcast = (bs == 0) || (B == 1);
    uint64_t dims[3] = {(uint64_t)rowlen, (uint64_t)L, (uint64_t)(bcast ? 1 : B)};
    uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)(bcast ? (uint64_t)ld * 2 * L : (uint64_t)bs * 2)};
    uint32_t box[3] = {AF_D, AF_T, 1};
    return make_tmap_bf16(tm, base, 3, dims, str, box, 64);
}

// floats of workspace a query-group launch needs (0: the call runs as one group and needs none)
static long long attn_bwd_fused_ws_floats(const fod_attn_args* a) {
    const int nqt = (a->Lq + AF_T - 1) / AF_T;
    if (a->q2 || nqt <= FusedBwdSmem<1>::kMaxQT) return 0;
    return 2ll * a->B * a->H * a->Lk * AF_D;
}

bool attn_bwd_fused_supported(const fod_attn_bwd_args* g) {
    const fod_attn_args* a = &g->fwd;
    const int nqt = (a->Lq + AF_T - 1) / AF_T;
    if (nqt <= (a->q2 ? FusedBwdSmem<2>::kMaxQT : FusedBwdSmem<1>::kMaxQT)) return true;
    // longer one-part sequences: query groups with dK / dV reduced through the caller's fp32 workspace
    return !a->q2 && g->ws && g->ws_floats >= attn_bwd_fused_ws_floats(a) && !(reinterpret_cast<uintptr_t>(g->ws) & 15);
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
    // gradient outputs (always carry a batch dimension; dq1 / dk1 / dv may be column slices of a fused tensor)
    const long long ld_dq = g->ld_dq ? g->ld_dq : rowlen, ld_dk = g->ld_dk ? g->ld_dk : rowlen;
    auto out_map = [&](CUtensorMap* tm, const void* base, int Lrows, long long ld) {
        uint64_t dims[3] = {(uint64_t)rowlen, (uint64_t)Lrows, (uint64_t)a->B};
        uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)ld * 2 * Lrows};
        uint32_t box[3] = {AF_D, AF_T, 1};
        return make_tmap_bf16(tm, base, 3, dims, str, box, 64);
    };
    CUtensorMap tdq1, tdq2, tdk1, tdk2, tdv;
    if (out_map(&tdq1, g->dq1, a->Lq, ld_dq) || out_map(&tdk1, g->dk1, a->Lk, ld_dk) || out_map(&tdv, g->dv, a->Lk, ld_dk))
        return fod_fail(FOD_E_TMAP, "attn_bwd: gradient outputs (16-byte aligned bases and row strides needed)");
    if (two) {
        if (out_map(&tdq2, g->dq2, a->Lq, rowlen) || out_map(&tdk2, g->dk2, a->Lk, rowlen))
            return fod_fail(FOD_E_TMAP, "attn_bwd: dq2 / dk2");
    } else {
        tdq2 = tdq1;
        tdk2 = tdk1;
    }
    AttnBwdFusedParams p;
    p.lse = a->lse;
    p.delta = g->delta;
    p.out = reinterpret_cast<const __nv_bfloat16*>(a->out);
    p.dout = reinterpret_cast<const __nv_bfloat16*>(g->dout);
    p.ldo = a->ldo;
    p.bso = a->bso;
    p.lddo = g->lddo;
    p.bsdo = a->B > 1 ? g->bsdo : 0;
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
    p.drop = {nullptr, 0u, 0u, 1.f};
    p.keep_bits = nullptr;
    p.keep_pitch = 0;
    const bool drop = a->drop.rng_state && a->drop.p > 0.f;
    if (drop) {
        if (!(a->drop.p < 1.f)) return fod_fail(FOD_E_ARG, "attn_bwd: dropout p must be in [0, 1)");
        p.drop.state = reinterpret_cast<const unsigned long long*>(a->drop.rng_state);
        p.drop.call = (unsigned int)a->drop.call;
        p.drop.thresh = (unsigned int)(a->drop.p * 65536.f + 0.5f);
        p.drop.scale = 1.f / (1.f - a->drop.p);
        if (a->keep_bits) {
            if ((a->keep_pitch % 16) || a->keep_pitch < 16LL * ((a->Lk + AF_T - 1) / AF_T) || (reinterpret_cast<uintptr_t>(a->keep_bits) & 15))
                return fod_fail(FOD_E_ARG, "attn_bwd: keep_bits needs a 16-byte aligned base and pitch >= 16 * ceil(Lk / 128), pitch % 16 == 0");
            p.keep_bits = reinterpret_cast<const uint8_t*>(a->keep_bits);
            p.keep_pitch = a->keep_pitch;
        }
    }
    const bool bits = p.keep_bits != nullptr;
    const int nqt_all = (a->Lq + AF_T - 1) / AF_T;
    const int groups = two ? 1 : (nqt_all + FusedBwdSmem<1>::kMaxQT - 1) / FusedBwdSmem<1>::kMaxQT;
    p.ws_dk = p.ws_dv = nullptr;
    if (groups > 1) {
        const long long half = (long long)a->B * a->H * a->Lk * AF_D;
        p.ws_dk = g->ws;
        p.ws_dv = g->ws + half;
        attn_ws_zero_kernel<<<148 * 4, 256, 0, st>>>(reinterpret_cast<float4*>(g->ws), 2 * half / 4);
    }
    dim3 grid(a->H, a->B, groups);
    auto launch = [&](auto kernel, int smem_bytes, bool& attr_done) {
        if (!attr_done) {
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
            attr_done = true;
        }
        fod_launch(kernel, grid, dim3(AF_THREADS), (size_t)smem_bytes, st, tq1, tk1, tq2, tk2, tv, tdo, tdq1, tdq2, tdk1, tdk2, tdv, p);
    };
    static bool attr[6] = {false, false, false, false, false, false};
    if (two) {
        if (drop && bits) launch(attn_bwd_fused_kernel<2, 2>, FusedBwdSmem<2>::kTotal, attr[5]);
        else if (drop) launch(attn_bwd_fused_kernel<2, 1>, FusedBwdSmem<2>::kTotal, attr[3]);
        else launch(attn_bwd_fused_kernel<2, 0>, FusedBwdSmem<2>::kTotal, attr[2]);
    } else {
        if (drop && bits) launch(attn_bwd_fused_kernel<1, 2>, FusedBwdSmem<1>::kTotal, attr[4]);
        else if (drop) launch(attn_bwd_fused_kernel<1, 1>, FusedBwdSmem<1>::kTotal, attr[1]);
        else launch(attn_bwd_fused_kernel<1, 0>, FusedBwdSmem<1>::kTotal, attr[0]);
    }
    int rc = fod_check_launch("attn_bwd_fused_kernel");
    if (rc || groups == 1) return rc;
    attn_ws_to_bf16_kernel<<<148 * 8, 256, 0, st>>>(p.ws_dk, p.ws_dv, p.dk1, p.dv, p.ld_dk, a->B, a->H, a->Lk);
    return fod_check_launch("attn_ws_to_bf16_kernel");
}

}  // namespace fod

#if FOD_AF_TRACE
extern "C" int fod_debug_af_trace(long long* host_out) {
    cudaDeviceSynchronize();
    return (int)cudaMemcpyFromSymbol(host_out, fod::g_af_trace, sizeof(long long) * 3 * 16 * 8);
}
#endif
