// Flash-attention forward for sm_100a: TMA-staged Q/K/V tiles, tcgen05.mma with TMEM accumulators,
// online softmax in registers.  Head dim is 32 per "part"; NPARTS = 2 realises the ConditionalDETR
// concatenated query/key ([content ‖ sine], head dim 64) as two accumulating MMAs into the same S tile.
//
//   warps 0..3 : softmax (one query row per thread; TMEM lane = row).  The whole 128-wide S row is pulled into
//                registers with four back-to-back tcgen05.ld (one wait), so S is read from TMEM once per tile.
//   warp 4     : TMA producer  (Q once, K/V double-buffered)
//   warp 5     : TMEM alloc + tcgen05.mma issuer:  S = Q·Kᵀ (M128 N128 K32·NPARTS),  O += P·V (M128 N32 K128) and
//                L += P·1 (same shape against a constant tile of ones): the row sums come from the tensor core, from
//                exactly the bf16 P that multiplies V.
//
// O and L ACCUMULATE IN TMEM across key tiles.  The exponent offset m_used is only raised when the running row
// maximum exceeds it by more than 2^8 (lazy rescaling): then the softmax warp multiplies its O/L row in TMEM by
// exp2(m_used - m_new).  In the steady state a softmax warp never waits for a P·V product; it reads O and L once.
//
// Per (batch, head, 128-query tile) CTA; TMEM (256 columns) admits 2 CTAs/SM, which overlap each other's phases.
// Bound: MUFU ex2 (one exp per 128 flop at head dim 32) and HBM/latency at S=350; see DESIGN.md.
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"
#include "dropout.cuh"

namespace fod {

constexpr int AT_BM = 128;  // queries per CTA
constexpr int AT_BN = 128;  // keys per tile
constexpr int AT_D = 32;    // head dim per part (and value head dim)
constexpr int AT_THREADS = 192;
constexpr int AT_TILE_BYTES = AT_BN * AT_D * 2;  // 8 KB, rows of 64 B, SWIZZLE_64B
constexpr int AT_P_BYTES = AT_BM * AT_BN * 2;    // 32 KB, two SWIZZLE_128B slabs of [128][64]
constexpr int AT_KV_STAGES = 2;
#ifndef FOD_ATTN_EARLY_S
#define FOD_ATTN_EARLY_S 1
#endif
#ifndef FOD_ATTN_DEBUG
#define FOD_ATTN_DEBUG 0
#endif
#if FOD_ATTN_DEBUG
// first few anomalies seen by the softmax warps: 16 floats per record
__device__ float g_attn_dbg[16 * 8];
__device__ int g_attn_dbg_n;
__device__ __noinline__ void attn_dbg_record(float kind, float b, float h, float q, float j, float nkt, float m_used,
                                             float mx, float l, float o0, float o1, float s0, float s1, float extra) {
    const int slot = atomicAdd(&g_attn_dbg_n, 1);
    if (slot >= 8) return;
    float* r = g_attn_dbg + slot * 16;
    r[0] = kind; r[1] = b; r[2] = h; r[3] = q; r[4] = j; r[5] = nkt; r[6] = m_used; r[7] = mx; r[8] = l; r[9] = o0;
    r[10] = o1; r[11] = s0; r[12] = s1; r[13] = extra; r[14] = (float)blockIdx.x; r[15] = (float)threadIdx.x;
}
#endif

template <int NPARTS>
struct AttnSmem {
    static constexpr int kQ = 0;
    static constexpr int kK = kQ + NPARTS * AT_TILE_BYTES;
    static constexpr int kV = kK + AT_KV_STAGES * NPARTS * AT_TILE_BYTES;
    static constexpr int kP = kV + AT_KV_STAGES * AT_TILE_BYTES;
    // one-part attention double-buffers P: the softmax warps of tile j only wait for P·V(j-2); with two parts the
    // second buffer would cost the second resident CTA
    static constexpr int kPBuf = NPARTS == 1 ? 2 : 1;
    static constexpr int kOnes = kP + kPBuf * AT_P_BYTES;            // [128 keys][32] bf16 ones: B operand of the row-sum MMA
    static constexpr int kBars = kOnes + AT_TILE_BYTES;
    // the dynamic window starts 1024-aligned (checked in the kernel): no alignment slack, which is what lets two
    // 112 KB CTAs share an SM
    static constexpr int kTotal = kBars + 128;
};

struct AttnParams {
    __nv_bfloat16* out;
    float* lse;
    long long ldo, bso;
    int B, H, Lq, Lk;
    int bm_q1, bm_k1, bm_q2, bm_k2, bm_v;  // 1 if the tensor has a batch dimension, 0 if broadcast
    float scale_log2;                      // scale * log2(e)
    float scale;
    DropoutCfg drop;                       // DROP kernels: dropout on the softmax weights (attention.py:370)
    const uint8_t* keep_bits;              // DROP == 2: the keep decisions, precomputed (fod_attn_keep_bits)
    long long keep_pitch;                  // bytes per (b, h, q) row: 16 per 128-key tile
};

// DROP: train-mode dropout on P.  out = dropout(P)·V / sum(P): the row sums must be those of the UNdropped weights, so
// they are accumulated in a register of the row's thread instead of by the [V | ones] MMA (whose L block then holds the
// dropped sums and is ignored).  Mask element index: ((b*H + h)*Lq + q) * roundup8(Lk) + k.
template <int NPARTS, int DROP>  // DROP: 0 none, 1 masks drawn here (Philox), 2 masks read from keep_bits
__global__ void __launch_bounds__(AT_THREADS, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ1, const __grid_constant__ CUtensorMap tmK1,
                const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmK2,
                const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    using L = AttnSmem<NPARTS>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    if (smem_u32(smem_raw) & 1023u) __trap();  // the swizzled tiles need a 1024-byte aligned window (no static smem here)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBars);
    uint64_t* q_full = bars + 0;
    uint64_t* kv_full = bars + 1;   // [2]
    uint64_t* kv_empty = bars + 3;  // [2]
    uint64_t* s_full = bars + 5;
    uint64_t* p_full = bars + 6;
    // P·V(j) completion alternates between two barriers: a parity wait can only tell the latest phase from the one
    // before, and with double-buffered P the softmax warps look two tiles back
    auto pv_full = [&](int j) { return bars + ((j & 1) ? 9 : 7); };
    uint64_t* s_free = bars + 8;    // softmax warps have copied S(j) into registers: S(j+1) may overwrite it
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * AT_BM;
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int nkt = (p.Lk + AT_BN - 1) / AT_BN;

    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmQ1);
        tma_prefetch_desc(&tmK1);
        tma_prefetch_desc(&tmV);
        if (NPARTS == 2) {
            tma_prefetch_desc(&tmQ2);
            tma_prefetch_desc(&tmK2);
        }
        mbar_init(q_full, 1);
        for (int s = 0; s < AT_KV_STAGES; ++s) {
            mbar_init(kv_full + s, 1);
            mbar_init(kv_empty + s, 1);
        }
        mbar_init(s_full, 1);
        mbar_init(p_full, 128);
        mbar_init(pv_full(0), 1);
        mbar_init(pv_full(1), 1);
        mbar_init(s_free, 128);
        fence_barrier_init();
    }
    if (warp == 5) tmem_alloc<256>(tmem_slot);
    if (warp < 4) {  // constant tile of ones (every element equal, so the swizzle pattern is irrelevant)
        uint4* ones = reinterpret_cast<uint4*>(smem + L::kOnes);
        const uint32_t one2 = 0x3F803F80u;  // bf16x2 (1.0, 1.0)
        for (int i = threadIdx.x; i < AT_TILE_BYTES / 16; i += 128) ones[i] = make_uint4(one2, one2, one2, one2);
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();  // the next kernel's prologue may overlap this grid (it still waits for our completion)
    pdl_wait();               // ours overlapped the previous kernel up to here; global memory is touched below only
    const uint32_t tmem_S = tmem_base;
    const uint32_t tmem_O = tmem_base + 128;  // 32 columns: O accumulator
    const uint32_t tmem_L = tmem_base + 160;  // 32 columns, all equal: row sums of P

    if (warp == 4) {
        if (lane == 0) {
            mbar_arrive_expect_tx(q_full, NPARTS * AT_TILE_BYTES);
            tma_load_3d(smem + L::kQ, &tmQ1, q_full, h * AT_D, q0, b * p.bm_q1);
            if (NPARTS == 2) tma_load_3d(smem + L::kQ + AT_TILE_BYTES, &tmQ2, q_full, h * AT_D, q0, b * p.bm_q2);
            for (int j = 0; j < nkt; ++j) {
                const int s = j % AT_KV_STAGES;
                const uint32_t ph = (j / AT_KV_STAGES) & 1;
                mbar_wait(kv_empty + s, ph ^ 1);
                mbar_arrive_expect_tx(kv_full + s, (NPARTS + 1) * AT_TILE_BYTES);
                uint8_t* sk = smem + L::kK + s * NPARTS * AT_TILE_BYTES;
                tma_load_3d(sk, &tmK1, kv_full + s, h * AT_D, j * AT_BN, b * p.bm_k1);
                if (NPARTS == 2) tma_load_3d(sk + AT_TILE_BYTES, &tmK2, kv_full + s, h * AT_D, j * AT_BN, b * p.bm_k2);
                tma_load_3d(smem + L::kV + s * AT_TILE_BYTES, &tmV, kv_full + s, h * AT_D, j * AT_BN, b * p.bm_v);
            }
        }
    } else if (warp == 5) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN, 0, 0);
            // O += P·V and L += P·1 are ONE instruction per 16 keys: B = [V | ones] with N = 64 (two MN-major chunks of 32
            // columns, the second one LBO bytes after the first), D = the adjacent O and L column blocks.  These skinny
            // MMAs are bound by the fetch of their A operand (the 128x16 P slice, ~65 clk whatever N is): issuing the
            // row-sum product separately doubled that cost.
            constexpr uint32_t idesc_pv = umma_idesc_bf16(AT_BM, 2 * AT_D, 0, 1);  // B is MN-major
            // stage-0 descriptors, built once and stepped with one add per instruction
            const uint64_t dQ0 = umma_smem_desc(smem_u32(smem + L::kQ), 16, 512, 4);
            const uint64_t dK0 = umma_smem_desc(smem_u32(smem + L::kK), 16, 512, 4);
            const uint64_t dP0 = umma_smem_desc(smem_u32(smem + L::kP), 16, 1024, 2);
            const uint32_t v0_addr = smem_u32(smem + L::kV), ones_addr = smem_u32(smem + L::kOnes);
            auto issue_S = [&](int j) {
                const int s = j % AT_KV_STAGES;
                mbar_wait(kv_full + s, (j / AT_KV_STAGES) & 1);
                tc_fence_after();
                const uint64_t dk = umma_desc_adv(dK0, s * NPARTS * AT_TILE_BYTES);
#pragma unroll
                for (int part = 0; part < NPARTS; ++part) {
#pragma unroll
                    for (int k = 0; k < AT_D / 16; ++k)
                        umma_bf16_ss(tmem_S, umma_desc_adv(dQ0, part * AT_TILE_BYTES + k * 32),
                                     umma_desc_adv(dk, part * AT_TILE_BYTES + k * 32), idesc_s, (part | k) ? 1u : 0u);
                }
                umma_commit(s_full);
            };
            mbar_wait(q_full, 0);
            issue_S(0);
            for (int j = 0; j < nkt; ++j) {
                const int s = j % AT_KV_STAGES;
                // S(j) lives in the softmax warps' registers as soon as they have loaded it: the next score tile is
                // computed while they exponentiate this one
#if FOD_ATTN_EARLY_S
                if (j + 1 < nkt) {
                    mbar_wait(s_free, j & 1);
                    tc_fence_after();
                    issue_S(j + 1);
                }
#endif
                mbar_wait(p_full, j & 1);
                tc_fence_after();
                const uint32_t v_addr = v0_addr + s * AT_TILE_BYTES;
                // [V | ones]: 16 keys per instruction = 16 rows of 64 B (+1024 B in both chunks); chunk 1 (the ones tile,
                // every element equal, so row offsets inside it do not matter) lies ones_addr - v_addr bytes after chunk 0
                const uint64_t dv = umma_smem_desc(v_addr, ones_addr - v_addr, 512, 4);
                const uint64_t dp = umma_desc_adv(dP0, (j % L::kPBuf) * AT_P_BYTES);
#pragma unroll
                for (int kk = 0; kk < AT_BN / 16; ++kk)
                    umma_bf16_ss(tmem_O, umma_desc_adv(dp, (kk >> 2) * (AT_BM * 128) + (kk & 3) * 32), umma_desc_adv(dv, kk * 1024),
                                 idesc_pv, (j | kk) ? 1u : 0u);  // O and L accumulate across key tiles
                umma_commit(pv_full(j));
                umma_commit(kv_empty + s);
#if !FOD_ATTN_EARLY_S
                if (j + 1 < nkt) issue_S(j + 1);
#endif
            }
        }
    } else {
        // ------------------------------------------------ softmax warps
        const int row = warp * 32 + lane;  // warp == TMEM lane quarter
        const uint32_t lane_off = uint32_t(warp * 32) << 16;
        uint8_t* sP = smem + L::kP;
        float m_used = -INFINITY;  // exponent offset in use for this row (raw score units)
        float l_reg = 0.f;         // DROP: running sum of the undropped weights of this row
        PhiloxKey key = {0u, 0u, 0u, 0u};
        unsigned long long g_row = 0;
        const uint4* bits_row = nullptr;
        if (DROP == 1) {
            key = dropout_key(p.drop);
            const unsigned long long lkp = (unsigned long long)((p.Lk + 7) & ~7);
            g_row = ((((unsigned long long)b * p.H + h) * p.Lq + (unsigned long long)min(q0 + row, p.Lq - 1)) * lkp) >> 3;
        }
        if (DROP == 2)
            bits_row = reinterpret_cast<const uint4*>(
                p.keep_bits + (((long long)b * p.H + h) * p.Lq + min(q0 + row, p.Lq - 1)) * p.keep_pitch);
        uint32_t v[4][32];
        for (int j = 0; j < nkt; ++j) {
            uint4 kb = make_uint4(0u, 0u, 0u, 0u);
            if (DROP == 2) kb = __ldg(bits_row + j);  // 128 keep bits of this row and key tile (in flight during the wait)
            mbar_wait(s_full, j & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, v[c]);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(s_free);
            const int kvalid = p.Lk - j * AT_BN;  // >= 1
            if (kvalid < AT_BN) {                 // ragged last tile (warp-uniform): masked keys score -inf
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i >= kvalid) v[c][i] = 0xff800000u;
            }
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;  // four independent chains
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                mx0 = fmaxf(mx0, __uint_as_float(v[0][i]));
                mx1 = fmaxf(mx1, __uint_as_float(v[1][i]));
                mx2 = fmaxf(mx2, __uint_as_float(v[2][i]));
                mx3 = fmaxf(mx3, __uint_as_float(v[3][i]));
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
#if FOD_ATTN_DEBUG
            {
                bool anynan = false;
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i) anynan |= (__uint_as_float(v[c][i]) != __uint_as_float(v[c][i]));
                if (anynan || !(fabsf(mx) < 3.0e38f))
                    attn_dbg_record(1.f, b, h, q0 + row, j, nkt, m_used, mx, 0.f, 0.f, 0.f, __uint_as_float(v[0][0]),
                                    __uint_as_float(v[3][31]), anynan ? 1.f : 0.f);
            }
#endif
            // lazy rescaling: keep the offset unless the row maximum outgrew it by more than 2^8
            const bool grow = (mx - m_used) * p.scale_log2 > 8.f;  // true on the first tile (m_used = -inf)
            if (__any_sync(0xffffffffu, grow)) {
                if (j > 0) {
                    // O / L of this row hold sums taken with the old offset: bring them to the new one.  Every P·V
                    // issued so far must have landed (the next one is only issued after this warp's p_full arrive).
                    mbar_wait(pv_full(j - 1), ((j - 1) >> 1) & 1);
                    tc_fence_after();
                    const float alpha = grow ? fast_exp2((m_used - mx) * p.scale_log2) : 1.f;
                    uint32_t o[32];
                    uint32_t l1[1];
                    tmem_ld_32x32b_x32(tmem_O + lane_off, o);
                    tmem_ld_32x32b_x1(tmem_L + lane_off, l1);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    l1[0] = __float_as_uint(__uint_as_float(l1[0]) * alpha);
                    tmem_st_32x32b_x32(tmem_O + lane_off, o);
                    tmem_st_32x32b_x1(tmem_L + lane_off, l1);
                    tmem_st_wait();
                    if (DROP) l_reg *= alpha;
                }
                if (grow) m_used = mx;
            }
            const float moff = m_used * p.scale_log2;
            // the P buffer about to be overwritten must have been consumed by its P·V (two tiles ago when double-buffered)
            if (j >= L::kPBuf) mbar_wait(pv_full(j - L::kPBuf), ((j - L::kPBuf) >> 1) & 1);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                // 32 bf16 = four 16-byte chunks: logical chunk index within the 64-wide slab
                uint8_t* slab = sP + (j % L::kPBuf) * AT_P_BYTES + (c >> 1) * (AT_BM * 128);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float e[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) e[t] = fast_exp2(fmaf(__uint_as_float(v[c][q * 8 + t]), p.scale_log2, -moff));
                    if (DROP) {
                        l_reg += ((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[7]));
                        // dropped weights are zeroed here; the 1/(1-p) of the kept ones is folded into the final 1/l
                        if (DROP == 1) {
                            const Keep8 keep = dropout_draw8(key, g_row + (unsigned long long)(j * (AT_BN / 8) + c * 4 + q), p.drop.thresh);
#pragma unroll
                            for (int t = 0; t < 8; ++t) e[t] = keep[t] ? e[t] : 0.f;
                        } else {
                            const uint32_t kw = (c == 0 ? kb.x : c == 1 ? kb.y : c == 2 ? kb.z : kb.w) >> (8 * q);
#pragma unroll
                            for (int t = 0; t < 8; ++t) e[t] = (kw & (1u << t)) ? e[t] : 0.f;
                        }
                    }
                    uint4 u;
                    u.x = pack_bf16x2(e[0], e[1]);
                    u.y = pack_bf16x2(e[2], e[3]);
                    u.z = pack_bf16x2(e[4], e[5]);
                    u.w = pack_bf16x2(e[6], e[7]);
                    *reinterpret_cast<uint4*>(slab + swz_off<128>(row, (c & 1) * 4 + q)) = u;
                }
            }
            fence_proxy_async_smem();
            tc_fence_before();
            mbar_arrive(p_full);
        }
        // every P·V and P·1 has been accumulated: read the row once
        mbar_wait(pv_full(nkt - 1), ((nkt - 1) >> 1) & 1);
        tc_fence_after();
        uint32_t o[32];
        uint32_t l1[1];
        tmem_ld_32x32b_x32(tmem_O + lane_off, o);
        tmem_ld_32x32b_x1(tmem_L + lane_off, l1);
        tmem_ld_wait();
        const int q = q0 + row;
        if (q < p.Lq) {
            const float l_run = DROP ? l_reg : __uint_as_float(l1[0]);
            const float inv = (DROP ? p.drop.scale : 1.f) / l_run;
#if FOD_ATTN_DEBUG
            {
                bool bad = !(l_run >= 0.5f) || !(l_run < 3.0e38f) || !(fabsf(m_used) < 3.0e38f);
#pragma unroll
                for (int i = 0; i < 32; ++i) bad |= !(fabsf(__uint_as_float(o[i])) < 3.0e38f);
                if (bad)
                    attn_dbg_record(2.f, b, h, q, nkt, nkt, m_used, (float)p.Lk, l_run, __uint_as_float(o[0]), __uint_as_float(o[1]),
                                    __uint_as_float(o[30]), __uint_as_float(o[31]), (float)p.Lq);
            }
#endif
            __nv_bfloat16* dst = p.out + (long long)b * p.bso + (long long)q * p.ldo + h * AT_D;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
                uint4 u;
                u.x = pack_bf16x2(__uint_as_float(o[qq * 8 + 0]) * inv, __uint_as_float(o[qq * 8 + 1]) * inv);
                u.y = pack_bf16x2(__uint_as_float(o[qq * 8 + 2]) * inv, __uint_as_float(o[qq * 8 + 3]) * inv);
                u.z = pack_bf16x2(__uint_as_float(o[qq * 8 + 4]) * inv, __uint_as_float(o[qq * 8 + 5]) * inv);
                u.w = pack_bf16x2(__uint_as_float(o[qq * 8 + 6]) * inv, __uint_as_float(o[qq * 8 + 7]) * inv);
                reinterpret_cast<uint4*>(dst)[qq] = u;
            }
            if (p.lse) p.lse[((long long)b * p.H + h) * p.Lq + q] = m_used * p.scale + logf(l_run);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 5) {
        tc_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

static int make_seq_tmap(CUtensorMap* tm, const void* base, int rowlen, int L, int B, long long ld, long long bs) {
    const bool bcast = (bs == 0) || (B == 1);
    uint64_t dims[3] = {(uint64_t)rowlen, (uint64_t)L, (uint64_t)(bcast ? 1 : B)};
    uint64_t str[2] = {(uint64_t)ld * 2, (uint64_t)(bcast ? (uint64_t)ld * 2 * L : (uint64_t)bs * 2)};
    uint32_t box[3] = {AT_D, AT_BN, 1};
    return make_tmap_bf16(tm, base, 3, dims, str, box, 64);
}

}  // namespace fod

#if FOD_ATTN_DEBUG
extern "C" int fod_attn_debug_read(float* host_records, int* host_count) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(host_records, fod::g_attn_dbg, sizeof(float) * 16 * 8);
    cudaMemcpyFromSymbol(host_count, fod::g_attn_dbg_n, sizeof(int));
    return 0;
}
#endif

extern "C" int fod_attn_fwd(const fod_attn_args* a, void* stream) {
    using namespace fod;
    if (!a || !a->q1 || !a->k1 || !a->v || !a->out) return fod_fail(FOD_E_ARG, "attn: null pointer");
    if (a->B <= 0 || a->H <= 0 || a->Lq <= 0 || a->Lk <= 0) return fod_fail(FOD_E_ARG, "attn: empty shape");
    if ((a->q2 == nullptr) != (a->k2 == nullptr)) return fod_fail(FOD_E_ARG, "attn: q2/k2 must come together");
    const int two = a->q2 != nullptr;
    const long long lds[6] = {a->ldq1, a->ldk1, a->ldv, a->ldo, two ? a->ldq2 : 8, two ? a->ldk2 : 8};
    for (int i = 0; i < 6; ++i)
        if (lds[i] % 8) return fod_fail(FOD_E_ARG, "attn: row strides must be multiples of 8 elements");
    const long long bss[6] = {a->bsq1, a->bsk1, a->bsv, a->bso, two ? a->bsq2 : 0, two ? a->bsk2 : 0};
    for (int i = 0; i < 6; ++i)
        if (bss[i] % 8) return fod_fail(FOD_E_ARG, "attn: batch strides must be multiples of 8 elements");
    const int rowlen = a->H * AT_D;
    CUtensorMap tq1, tk1, tq2, tk2, tv;
    if (make_seq_tmap(&tq1, a->q1, rowlen, a->Lq, a->B, a->ldq1, a->bsq1)) return fod_fail(FOD_E_TMAP, "attn: tmap q1");
    if (make_seq_tmap(&tk1, a->k1, rowlen, a->Lk, a->B, a->ldk1, a->bsk1)) return fod_fail(FOD_E_TMAP, "attn: tmap k1");
    if (make_seq_tmap(&tv, a->v, rowlen, a->Lk, a->B, a->ldv, a->bsv)) return fod_fail(FOD_E_TMAP, "attn: tmap v");
    if (two) {
        if (make_seq_tmap(&tq2, a->q2, rowlen, a->Lq, a->B, a->ldq2, a->bsq2)) return fod_fail(FOD_E_TMAP, "attn: tmap q2");
        if (make_seq_tmap(&tk2, a->k2, rowlen, a->Lk, a->B, a->ldk2, a->bsk2)) return fod_fail(FOD_E_TMAP, "attn: tmap k2");
    } else {
        tq2 = tq1;
        tk2 = tk1;
    }
    AttnParams p;
    p.out = reinterpret_cast<__nv_bfloat16*>(a->out);
    p.lse = a->lse;
    p.ldo = a->ldo;
    p.bso = a->bso;
    p.B = a->B;
    p.H = a->H;
    p.Lq = a->Lq;
    p.Lk = a->Lk;
    auto has_b = [&](long long bs) { return (bs != 0 && a->B > 1) ? 1 : 0; };
    p.bm_q1 = has_b(a->bsq1);
    p.bm_k1 = has_b(a->bsk1);
    p.bm_q2 = two ? has_b(a->bsq2) : 0;
    p.bm_k2 = two ? has_b(a->bsk2) : 0;
    p.bm_v = has_b(a->bsv);
    p.scale = a->scale;
    p.scale_log2 = a->scale * 1.4426950408889634f;
    p.drop = {nullptr, 0u, 0u, 1.f};
    p.keep_bits = nullptr;
    p.keep_pitch = 0;
    const bool drop = a->drop.rng_state && a->drop.p > 0.f;
    if (drop) {
        if (!(a->drop.p < 1.f)) return fod_fail(FOD_E_ARG, "attn: dropout p must be in [0, 1)");
        p.drop.state = reinterpret_cast<const unsigned long long*>(a->drop.rng_state);
        p.drop.call = (unsigned int)a->drop.call;
        p.drop.thresh = (unsigned int)(a->drop.p * 65536.f + 0.5f);
        p.drop.scale = 1.f / (1.f - a->drop.p);
        if (a->keep_bits) {
            if ((a->keep_pitch % 16) || a->keep_pitch < 16LL * ((a->Lk + AT_BN - 1) / AT_BN) || (reinterpret_cast<uintptr_t>(a->keep_bits) & 15))
                return fod_fail(FOD_E_ARG, "attn: keep_bits needs a 16-byte aligned base and pitch >= 16 * ceil(Lk / 128), pitch % 16 == 0");
            p.keep_bits = reinterpret_cast<const uint8_t*>(a->keep_bits);
            p.keep_pitch = a->keep_pitch;
        }
    }
    const bool bits = p.keep_bits != nullptr;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    dim3 grid((a->Lq + AT_BM - 1) / AT_BM, a->H, a->B);
    auto launch = [&](auto kernel, int smem_bytes, bool& attr_done) {
        if (!attr_done) {
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
            attr_done = true;
        }
        fod_launch(kernel, grid, dim3(AT_THREADS), (size_t)smem_bytes, st, tq1, tk1, tq2, tk2, tv, p);
    };
    static bool attr[6] = {false, false, false, false, false, false};
    if (two) {
        if (drop && bits) launch(attn_fwd_kernel<2, 2>, AttnSmem<2>::kTotal, attr[5]);
        else if (drop) launch(attn_fwd_kernel<2, 1>, AttnSmem<2>::kTotal, attr[3]);
        else launch(attn_fwd_kernel<2, 0>, AttnSmem<2>::kTotal, attr[2]);
    } else {
        if (drop && bits) launch(attn_fwd_kernel<1, 2>, AttnSmem<1>::kTotal, attr[4]);
        else if (drop) launch(attn_fwd_kernel<1, 1>, AttnSmem<1>::kTotal, attr[1]);
        else launch(attn_fwd_kernel<1, 0>, AttnSmem<1>::kTotal, attr[0]);
    }
    return fod_check_launch("attn_fwd_kernel");
}


// ---------------------------------------------------------------------------------------------------------------------
// Keep decisions of the softmax-weight dropout for one attention call, drawn once (forward and backward read them):
// word w of a row covers keys [32 w, 32 w + 32) = Philox groups 4 w .. 4 w + 3 of that row, byte i of the word = group
// 4 w + i, bit t = element t of the group — the lane order of dropout_keep8 / Keep8 (csrc/dropout.cuh).
namespace fod {
__global__ void __launch_bounds__(256)
attn_keep_bits_kernel(DropoutCfg d, long long rows, int lkp8, int words, uint32_t* __restrict__ out) {
    const PhiloxKey key = dropout_key(d);
    const long long total = rows * words;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long r = i / words;
        const int w = (int)(i - r * words);
        uint32_t m = 0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const int grp = 4 * w + g;
            if (grp < lkp8) m |= dropout_keep8(key, (unsigned long long)r * lkp8 + grp, d.thresh) << (8 * g);
        }
        out[i] = m;
    }
}
}  // namespace fod

extern "C" int fod_attn_keep_bits(const void* rng_state, int32_t call, float p, int64_t rows, int32_t Lk, void* out,
                                  int64_t pitch, void* stream) {
    using namespace fod;
    if (!rng_state || !out || rows <= 0 || Lk <= 0) return fod_fail(FOD_E_ARG, "attn_keep_bits: null/empty");
    if (!(p > 0.f) || !(p < 1.f)) return fod_fail(FOD_E_ARG, "attn_keep_bits: p must be in (0, 1)");
    if ((pitch % 16) || pitch < 16LL * ((Lk + 127) / 128) || (reinterpret_cast<uintptr_t>(out) & 15))
        return fod_fail(FOD_E_ARG, "attn_keep_bits: 16-byte aligned out, pitch % 16 == 0, pitch >= 16 * ceil(Lk / 128)");
    DropoutCfg d;
    d.state = reinterpret_cast<const unsigned long long*>(rng_state);
    d.call = (unsigned int)call;
    d.thresh = (unsigned int)(p * 65536.f + 0.5f);
    d.scale = 1.f / (1.f - p);
    const int words = (int)(pitch / 4);
    const long long total = rows * words;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    attn_keep_bits_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(d, rows, (Lk + 7) >> 3, words,
                                                                                          reinterpret_cast<uint32_t*>(out));
    return fod_check_launch("attn_keep_bits_kernel");
}
