// ResNet stem: 7x7 / stride 2 / pad 3 convolution, 3 -> 64 channels, with the frozen BatchNorm folded in and ReLU
// (ConditionalDETR/models/backbone.py:88-96 -> torchvision resnet conv1 + bn1 + relu; frozen, forward only:
// paper.py:102-109).  cuDNN runs this shape on an sm_80 implicit-GEMM kernel (0.82 ms for 24 x 448x800 frames, the largest
// single launch of the training step); with 3 input channels neither TMA im2col nor channel-vectorised loads apply.
//
// Implicit GEMM on tcgen05, A built by the threads:
//   tile = 8 x 16 output pixels (M = 128), N = 64 output channels, K = 7 kernel rows x 24 (= 1 + 7*3 + 2) = 168 -> 192.
//   For one kernel row the 7x3 input values of an output pixel are CONTIGUOUS in the NHWC image (42 bytes); starting one
//   element early makes the run 4-byte aligned, and the 3 surplus elements meet zero weights.  The input patch of a tile
//   (21 rows x 256 bytes) is staged with 16-byte cp.async (zero-filled outside the image; double-buffered across
//   tiles), each thread assembles its pixel's K row into the SWIZZLE_128B operand tile (84 ld.shared.b32, 21
//   st.shared.v4), one thread issues 12 MMAs (128x64x16), and the epilogue reads the accumulator row from TMEM, adds the
//   folded shift, applies ReLU and writes the pixel's 64 channels (128 contiguous bytes, NHWC).
// Persistent CTAs, two per SM, so one CTA's operand assembly overlaps the other's MMA + epilogue.
// HBM-bound in the limit: reads the image once (6 B/pixel), writes 128 B per output pixel (275 MB for the bench step).
#include <cstdlib>
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "tmap.h"

namespace fod {

constexpr int ST_TH = 8, ST_TW = 16;            // output tile
constexpr int ST_PROWS = 2 * ST_TH + 5;         // 21 input rows
constexpr int ST_PROW_BYTES = 256;              // 16 chunks: input bytes [192*tx - 32, 192*tx + 224) of an image row
constexpr int ST_KG = 24;                       // K elements per kernel row
constexpr int ST_K = 192;                       // 7 * 24 = 168, padded to 3 slabs of 64
static_assert(7 * ST_KG <= ST_K && ST_KG * 2 == 48, "one kernel row = three 16-byte chunks");
constexpr int ST_A_BYTES = 3 * 128 * 128;       // 3 slabs [128 rows][128 B]
constexpr int ST_B_BYTES = 3 * 64 * 128;        // 3 slabs [64 rows][128 B]
constexpr int ST_PATCH_BYTES = ST_PROWS * ST_PROW_BYTES;
constexpr int ST_SMEM = ST_A_BYTES + ST_B_BYTES + 2 * ST_PATCH_BYTES + 64 * 4 + 64 + 1024;

struct StemParams {
    const __nv_bfloat16* x;   // (N, H, W, 3)
    const __nv_bfloat16* wk;  // (64, 192): [o][kh*24 + 1 + kw*3 + c] = W[o][c][kh][kw] * s[o], zeros elsewhere
    const float* bias;        // (64) folded shift
    __nv_bfloat16* y;         // (N, OH, OW, 64)
    int N, H, W, OH, OW, tiles_x, tiles_y, tiles;
};

__device__ __forceinline__ void cp_async_16_zfill(uint32_t dst, const void* src, bool valid) {
    const int n = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}

__device__ __forceinline__ void stem_issue_patch(const StemParams& p, int tile, uint32_t dst_base) {
    const int tx = tile % p.tiles_x;
    const int t2 = tile / p.tiles_x;
    const int ty = t2 % p.tiles_y;
    const int n = t2 / p.tiles_y;
    const int iy0 = 2 * ty * ST_TH - 3;
    const long long row_bytes = (long long)p.W * 6;
    const long long col0 = 192LL * tx - 32;  // byte offset inside an image row, multiple of 16
    const char* img = reinterpret_cast<const char*>(p.x) + (long long)n * p.H * row_bytes;
    for (int c = threadIdx.x; c < ST_PROWS * 16; c += 128) {
        const int r = c >> 4, k = c & 15;
        const int iy = iy0 + r;
        const long long b = col0 + 16 * k;
        const bool valid = iy >= 0 && iy < p.H && b >= 0 && b + 16 <= row_bytes;
        const char* src = valid ? img + (long long)iy * row_bytes + b : reinterpret_cast<const char*>(p.x);
        cp_async_16_zfill(dst_base + r * ST_PROW_BYTES + k * 16, src, valid);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(128, 2)
stem_conv7x7s2_kernel(const StemParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = sA + ST_A_BYTES;
    uint8_t* sPatch = sB + ST_B_BYTES;
    float* sBias = reinterpret_cast<float*>(sPatch + 2 * ST_PATCH_BYTES);
    uint64_t* mma_bar = reinterpret_cast<uint64_t*>(sBias + 64);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(mma_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<64>(tmem_slot);
    // weights -> swizzled B tile (once); zero tail of every A row (k = 168..191 never changes); bias
    for (int c = tid; c < 64 * 24; c += 128) {
        const int o = c / 24, ch = c - o * 24;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.wk + (long long)o * ST_K) + ch);
        *reinterpret_cast<uint4*>(sB + (ch >> 3) * (64 * 128) + swz_off<128>(o, ch & 7)) = v;
    }
#pragma unroll
    for (int ch = 21; ch < 24; ++ch)
        *reinterpret_cast<uint4*>(sA + (ch >> 3) * (128 * 128) + swz_off<128>(tid, ch & 7)) = make_uint4(0, 0, 0, 0);
    if (tid < 64) sBias[tid] = p.bias[tid];
    int tile = blockIdx.x;
    if (tile < p.tiles) stem_issue_patch(p, tile, smem_u32(sPatch));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_acc = *tmem_slot;
    const uint32_t lane_off = uint32_t(warp * 32) << 16;
    constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
    const int py = tid >> 4, px = tid & 15;
    int buf = 0;
    uint32_t phase = 0;
    for (; tile < p.tiles; tile += gridDim.x) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();  // this tile's patch has landed; the previous tile's epilogue and MMA are done in every warp
        const int next = tile + gridDim.x;
        if (next < p.tiles) stem_issue_patch(p, next, smem_u32(sPatch + (buf ^ 1) * ST_PATCH_BYTES));
        // ---- this thread's pixel: 7 kernel rows x 48 bytes from the patch -> 21 chunks of the operand row
        const uint8_t* patch = sPatch + buf * ST_PATCH_BYTES;
#pragma unroll
        for (int kh = 0; kh < 7; ++kh) {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(patch + (2 * py + kh) * ST_PROW_BYTES + 12 + 12 * px);
            uint32_t w[12];
#pragma unroll
            for (int i = 0; i < 12; ++i) w[i] = src[i];
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const int ch = kh * 3 + q;
                *reinterpret_cast<uint4*>(sA + (ch >> 3) * (128 * 128) + swz_off<128>(tid, ch & 7)) =
                    make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
            }
        }
        fence_proxy_async_smem();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
#pragma unroll
            for (int kk = 0; kk < ST_K / 16; ++kk) {
                const uint64_t da = umma_smem_desc(a_addr + (kk >> 2) * (128 * 128) + (kk & 3) * 32, 16, 1024, 2);
                const uint64_t db = umma_smem_desc(b_addr + (kk >> 2) * (64 * 128) + (kk & 3) * 32, 16, 1024, 2);
                umma_bf16_ss(tmem_acc, da, db, idesc, kk ? 1u : 0u);
            }
            umma_commit(mma_bar);
        }
        mbar_wait(mma_bar, phase);
        phase ^= 1;
        tc_fence_after();
        // ---- epilogue: accumulator row -> +shift, ReLU -> 64 bf16 channels of this pixel
        uint32_t acc[2][32];
        tmem_ld_32x32b_x32(tmem_acc + lane_off, acc[0]);
        tmem_ld_32x32b_x32(tmem_acc + lane_off + 32, acc[1]);
        tmem_ld_wait();
        tc_fence_before();
        const int tx = tile % p.tiles_x;
        const int t2 = tile / p.tiles_x;
        const int ty = t2 % p.tiles_y;
        const int n = t2 / p.tiles_y;
        const int oy = ty * ST_TH + py, ox = tx * ST_TW + px;
        if (oy < p.OH && ox < p.OW) {
            uint4* dst = reinterpret_cast<uint4*>(p.y + (((long long)n * p.OH + oy) * p.OW + ox) * 64);
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float f[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) f[t] = fmaxf(__uint_as_float(acc[h][q * 8 + t]) + sBias[h * 32 + q * 8 + t], 0.f);
                    uint4 u;
                    u.x = pack_bf16x2(f[0], f[1]);
                    u.y = pack_bf16x2(f[2], f[3]);
                    u.z = pack_bf16x2(f[4], f[5]);
                    u.w = pack_bf16x2(f[6], f[7]);
                    dst[h * 4 + q] = u;
                }
        }
        buf ^= 1;
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<64>(tmem_acc);
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// Second form: the operand tile comes straight from TMA (no thread-built im2col).  The image is stored NHWC with FOUR
// channels (the fourth zero) and explicit zero columns left and right: (N, H, W + 8, 4) bf16, pixel x of the image at
// padded column x + 3.  For one kernel row the 7 x 3 inputs of output pixel ox are then the 32 contiguous elements
// (8 pixels x 4 channels, 64 bytes; the 8th pixel and every 4th channel meet zero weights) that start at padded column
// 2 ox — 16 bytes further for every output pixel.  A tiled tensor map whose second dimension has a 16-byte stride under
// a 64-byte innermost extent describes exactly that overlapping window (csrc/tmaprobe.cu checks it on the device), and
// splitting the image rows into (parity, row / 2) turns the stride-2 row walk into a dense box:
//     dims [k = 32][ox][row parity = 2][row / 2][n], box [32][16][1][11][1] -> [11 half-rows][16 pixels][64 B], SWIZZLE_64B:
// any 8 consecutive half-rows of it are the K-major A operand of one kernel row as they land, and the kernel rows of one
// parity (1,3,5 / 0,2,4,6) are the SAME box read one half-row further down (two boxes = 22 KB per tile, not seven).
// Rows above / below the image and tiles hanging over the right edge are zero-filled by the TMA unit.
// K = 7 x 32 = 224: 14 MMAs (M128 N64 K16) per 8 x 16-pixel tile.
//   warp 0: TMA producer, the two parity boxes of a tile on one barrier, 5-tile ring
//   warp 1: TMEM allocator + MMA issuer, two accumulators (2 x 64 columns): tile i+1 is multiplied while tile i drains
//   warps 2-5: epilogue — accumulator row -> +shift, ReLU -> bf16 -> SWIZZLE_128B staging tile -> ONE TMA store of
//              the tile's [8][16][64] block (clipped at the image edges by the unit).
// Persistent, one CTA per SM.  Measured (64 frames 448x800): 243 us = 1 520 clk per tile against 730 clk of MMA issue — the
// 128 B per output pixel leave at 3.0 TB/s where a B200 sustains 3.9 TB/s for pure stores (scripts/bw_probe.py), which
// is why the POOL form below exists.
constexpr int S4_STAGES = 5;                     // ring of whole tiles
constexpr int S4_BOX_ROWS = 11;                  // half-rows per box: 8 output rows + 3 (odd kernel rows 0,2,4,6 share one box)
constexpr int S4_BOX_BYTES = S4_BOX_ROWS * 16 * 64;  // 11 KB: [11 half-rows][16 pixels][64 B]
constexpr int S4_A_BYTES = 2 * S4_BOX_BYTES;     // even-parity box (kernel rows 1,3,5) + odd-parity box (0,2,4,6)
constexpr int S4_B_BYTES = 7 * 64 * 64;          // weights: 7 kernel rows x [64 out][32 k]
constexpr int S4_OUT_BYTES = 128 * 128;          // staged output tile
constexpr int S4_THREADS = 192;
// RAW form: the tile is 16 output rows x 8 pixels and shared memory holds the image half-rows AS THEY ARE (256 B each:
// the 22 padded columns the 8 pixels touch, rounded up to the box), not their im2col expansion.  The expansion is done by
// the MMA's own operand addressing: with the un-swizzled K-major layout a matrix descriptor places element (m, k) at
//     start + (m / 8) * SBO + (m % 8) * 16 + (k / 8) * LBO + (k % 8) * 2   bytes,
// and LBO = 16, SBO = 256 (one half-row per 8-pixel group) make that start + y * 256 + x * 16 + k * 2: the overlapping
// 64-byte window of output pixel (y, x).  Two boxes of 19 half-rows per tile (9.5 KB instead of 22 KB, 38 row requests
// of 256 B for the TMA unit instead of 352 of 64 B, which is what bounded the expanded form at ~1 500 clk per tile).
constexpr int S4R_BOX_ROWS = 19;
constexpr int S4R_ROW_BYTES = 256;
constexpr int S4R_BOX_BYTES = S4R_BOX_ROWS * S4R_ROW_BYTES;
constexpr int S4R_A_BYTES = 2 * S4R_BOX_BYTES;

struct Stem4Params {
    const __nv_bfloat16* wk;  // (64, 224): [o][kh*32 + kw*4 + c] = W[o][c][kh][kw] * s[o] (kw < 7, c < 3), zeros elsewhere
    const float* bias;
    __nv_bfloat16* pooled;    // POOL: (N, PH, PW, 64) output of the fused 3x3 / stride 2 / pad 1 max-pool
    int tiles_x, tiles_y, tiles;
    int OH, OW, PH, PW;       // convolution / pooled output extents
};

// POOL form (RAW tiles only): the max-pool that follows the stem (torchvision resnet `maxpool`, kernel 3, stride 2,
// padding 1) is taken inside the epilogue, so the 64-channel convolution output — 128 B per pixel, 734 MB for the bench
// step, whose WRITE alone costs 190 us at the 3.9 TB/s a B200 sustains for pure stores (scripts/bw_probe.py) — never
// leaves the SM; only the pooled quarter of it does.  A tile of 16 x 8 convolution outputs that starts at an odd row and
// column (14 ty - 1, 6 tx - 1) contains every input of a 7 x 3 block of pooled pixels: tiles overlap by two rows / columns
// (21 pooled pixels from 128 convolution outputs, 1.5x the MMAs — still below what the store cost).  Convolution
// positions outside the image are skipped in the maximum (everything kept is >= 0 after the ReLU).
constexpr int S4P_PH = 7, S4P_PW = 3;

// POOL runs G (default 3) epilogue groups of four warps (tiles go round-robin, one TMEM accumulator and one staging tile
// per group): with one group the pooling pass made the epilogue the longest stage of a tile (830 us; 2 groups 399,
// 3 groups 375, 4 groups 377 — beyond three the shared-memory traffic of MMA operands + staging + pooling binds).
template <bool POOL, int G> constexpr int s4_threads() { return POOL ? 64 + 128 * G : S4_THREADS; }
template <bool RAW, bool POOL, int G> constexpr int s4_smem() {
    // (every form keeps the expanded form's stage region, RAW stages being less than half of it: with the weights and
    // staging tiles placed right behind the smaller RAW ring the un-pooled launch took 288 us instead of 243, reproducibly)
    return ((S4_STAGES * S4_A_BYTES + 1023) & ~1023) + S4_B_BYTES + (POOL ? G : 2) * S4_OUT_BYTES + 64 * 4 + 512 + 1024;
}

template <bool RAW, bool POOL, int G>
__global__ void __launch_bounds__(s4_threads<POOL, G>(), 1)
stem_conv7x7s2_p4_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY, const Stem4Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int A_BYTES = RAW ? S4R_A_BYTES : S4_A_BYTES;
    constexpr int BOX_BYTES = RAW ? S4R_BOX_BYTES : S4_BOX_BYTES;
    constexpr int TW = RAW ? 8 : 16, TH = RAW ? 16 : 8;   // output tile (pixels x rows), M = 128 either way
    static_assert(RAW || !POOL, "the fused max-pool works on the 16 x 8 tiles of the RAW form");
    // first convolution row / column of tile (ty, tx)
    constexpr int SY = POOL ? 2 * S4P_PH : TH, SX = POOL ? 2 * S4P_PW : TW, O0 = POOL ? -1 : 0;
    uint8_t* sA = smem;
    constexpr int NB = POOL ? G : 2;                      // TMEM accumulators (64 columns each) = staging tiles
    constexpr int TCOLS = NB <= 2 ? 128 : 256;
    static_assert(NB <= 4, "barrier block holds four accumulator barriers");
    // (1 KB-aligned: the SWIZZLE_128B staging tiles follow; the non-POOL forms keep the expanded form's stage region)
    uint8_t* sB = sA + ((S4_STAGES * S4_A_BYTES + 1023) & ~1023);
    uint8_t* sOut = sB + S4_B_BYTES;
    float* sBias = reinterpret_cast<float*>(sOut + NB * S4_OUT_BYTES);
    uint64_t* full = reinterpret_cast<uint64_t*>(sBias + 64);
    uint64_t* empty = full + S4_STAGES;
    uint64_t* acc_full = empty + S4_STAGES;
    uint64_t* acc_empty = acc_full + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 4);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmY);
        for (int s = 0; s < S4_STAGES; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, 1);
        }
        for (int b = 0; b < NB; ++b) {
            mbar_init(acc_full + b, 1);
            mbar_init(acc_empty + b, 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TCOLS>(tmem_slot);
    // weights -> 7 swizzled [64][64 B] tiles (once per CTA); folded shift
    for (int c = tid; c < 64 * 28; c += s4_threads<POOL, G>()) {
        const int o = c / 28, ch = c - o * 28;  // 28 chunks of 8 k per output channel
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.wk + (long long)o * 224) + ch);
        *reinterpret_cast<uint4*>(sB + (ch >> 2) * (64 * 64) + swz_off<64>(o, ch & 3)) = v;
    }
    if (tid < 64) sBias[tid] = p.bias[tid];
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
                const int tx = tile % p.tiles_x;
                const int t2 = tile / p.tiles_x;
                const int ty = t2 % p.tiles_y;
                const int n = t2 / p.tiles_y;
                mbar_wait(empty + s, ph ^ 1);
                mbar_arrive_expect_tx(full + s, A_BYTES);
                // input row 2 oy - 3 + kh = 2 (oy + off) + par: kh = 1,3,5 -> par 0, off -1,0,1; kh = 0,2,4,6 -> par 1, off -2..1
                if constexpr (RAW) {  // [element of the padded row][parity][row / 2][n]: 8 output pixels start at padded column 16 tx
                    const int r0 = ty * SY + O0, c0 = tx * SX + O0;   // (padded column of output column c0: 2 c0, 4 channels each)
                    tma_load_4d(sA + s * A_BYTES, &tmX, full + s, c0 * 8, 0, r0 - 1, n);
                    tma_load_4d(sA + s * A_BYTES + BOX_BYTES, &tmX, full + s, c0 * 8, 1, r0 - 2, n);
                } else {
                    tma_load_5d(sA + s * A_BYTES, &tmX, full + s, 0, tx * TW, 0, ty * TH - 1, n);
                    tma_load_5d(sA + s * A_BYTES + BOX_BYTES, &tmX, full + s, 0, tx * TW, 1, ty * TH - 2, n);
                }
                if (++s == S4_STAGES) {
                    s = 0;
                    ph ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
            // expanded form: [128][64 B] SWIZZLE_64B, 8-row groups 512 B apart; RAW: un-swizzled, LBO 16 / SBO 256 (see above)
            const uint64_t dA0 = RAW ? umma_smem_desc(smem_u32(sA), 16, S4R_ROW_BYTES, 0) : umma_smem_desc(smem_u32(sA), 16, 512, 4);
            const uint64_t dB0 = umma_smem_desc(smem_u32(sB), 16, 512, 4);
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
                const int buf = it % NB;
                mbar_wait(acc_empty + buf, ((it / NB) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * 64;
                mbar_wait(full + s, ph);
                tc_fence_after();
#pragma unroll
                for (int kh = 0; kh < 7; ++kh) {
                    // kernel row kh starts (kh >> 1) half-rows (1 KB each: 16 pixels x 64 B, a multiple of the 512-byte
                    // swizzle period) into the box of its parity
                    const uint64_t da = umma_desc_adv(dA0, s * A_BYTES + ((kh & 1) ? 0 : BOX_BYTES) + (kh >> 1) * (RAW ? S4R_ROW_BYTES : 1024));
                    const uint64_t db = umma_desc_adv(dB0, kh * (64 * 64));
                    umma_bf16_ss(d_tmem, da, db, idesc, kh ? 1u : 0u);
                    umma_bf16_ss(d_tmem, umma_desc_adv(da, 32), umma_desc_adv(db, 32), idesc, 1u);
                }
                umma_commit(empty + s);
                umma_commit(acc_full + buf);
                if (++s == S4_STAGES) {
                    s = 0;
                    ph ^= 1;
                }
            }
        }
    } else {
        const int quarter = warp & 3;                 // TMEM lane quarter of this warp
        const int row = quarter * 32 + lane;          // pixel (row / TW, row % TW) of the tile
        const uint32_t lane_off = uint32_t(quarter * 32) << 16;
        const bool leader = tid == 64;
        constexpr int GROUPS = POOL ? G : 1;
        const int group = (warp - 2) >> 2;            // 0 unless POOL
        const int gtid = tid - 64 - group * 128;
        const int bar_a = 1 + 2 * group, bar_b = 2 + 2 * group;
        // POOL: this thread's (pooled pixel, 16-byte channel chunk) items — the same in every tile — with the nine staged
        // addresses of each window precomputed
        uint32_t p_off[2][9];
        int p_rel[2], p_pr[2], p_pc[2];
        bool p_on[2];
        if constexpr (POOL) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int item = gtid + 128 * j, q = item & 7, pp = item >> 3;
                p_on[j] = item < S4P_PH * S4P_PW * 8;
                p_pr[j] = pp / S4P_PW;
                p_pc[j] = pp - p_pr[j] * S4P_PW;
#pragma unroll
                for (int d = 0; d < 9; ++d)
                    p_off[j][d] = p_on[j] ? swz_off<128>((2 * p_pr[j] + d / 3) * TW + 2 * p_pc[j] + d % 3, q) : 0u;
                p_rel[j] = (p_pr[j] * p.PW + p_pc[j]) * 64 + q * 8;
            }
        }
        for (int it = group, tile = blockIdx.x + group * gridDim.x; tile < p.tiles; tile += GROUPS * gridDim.x, it += GROUPS) {
            const int buf = it % NB;                  // (POOL: == group)
            mbar_wait(acc_full + buf, (it / NB) & 1);
            tc_fence_after();
            uint32_t acc[2][32];
            tmem_ld_32x32b_x32(tmem_base + buf * 64 + lane_off, acc[0]);
            tmem_ld_32x32b_x32(tmem_base + buf * 64 + lane_off + 32, acc[1]);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty + buf);  // the MMA warp may overwrite this accumulator
            // the staging tile `buf` was read by the store issued two tiles ago: the leader waited for it (below)
            asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
            uint8_t* so = sOut + buf * S4_OUT_BYTES;
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float f[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        // POOL: the ReLU is the zero the pooled maximum starts from (max and bf16 rounding are monotone:
                        // max_i bf16(relu(x_i)) == max(0, max_i bf16(x_i)) bit for bit)
                        const float x = __uint_as_float(acc[h][q * 8 + t]) + sBias[h * 32 + q * 8 + t];
                        f[t] = POOL ? x : fmaxf(x, 0.f);
                    }
                    uint4 u;
                    u.x = pack_bf16x2(f[0], f[1]);
                    u.y = pack_bf16x2(f[2], f[3]);
                    u.z = pack_bf16x2(f[4], f[5]);
                    u.w = pack_bf16x2(f[6], f[7]);
                    *reinterpret_cast<uint4*>(so + swz_off<128>(row, h * 4 + q)) = u;
                }
            if constexpr (!POOL) fence_proxy_async_smem();
            if constexpr (POOL) {
                asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");  // the staged tile is complete
                const int tx = tile % p.tiles_x;
                const int t2 = tile / p.tiles_x;
                const int ty = t2 % p.tiles_y;
                const int n = t2 / p.tiles_y;
                const int r0 = ty * SY + O0, c0 = tx * SX + O0;
                const __nv_bfloat162 zero = __floats2bfloat162_rn(0.f, 0.f);
                if (r0 >= 0 && r0 + 2 * S4P_PH < p.OH && c0 >= 0 && c0 + 2 * S4P_PW < p.OW && ty * S4P_PH + S4P_PH <= p.PH &&
                    tx * S4P_PW + S4P_PW <= p.PW) {  // interior tile (CTA-uniform): every window complete
                    __nv_bfloat16* blk = p.pooled + (((long long)n * p.PH + ty * S4P_PH) * p.PW + tx * S4P_PW) * 64;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        if (!p_on[j]) continue;
                        uint4 v[9];
#pragma unroll
                        for (int d = 0; d < 9; ++d) v[d] = *reinterpret_cast<const uint4*>(so + p_off[j][d]);
                        __nv_bfloat162* hb = reinterpret_cast<__nv_bfloat162*>(&v[0]);
#pragma unroll
                        for (int d = 1; d < 9; ++d) {
                            const __nv_bfloat162* hv = reinterpret_cast<const __nv_bfloat162*>(&v[d]);
#pragma unroll
                            for (int e = 0; e < 4; ++e) hb[e] = __hmax2(hb[e], hv[e]);
                        }
#pragma unroll
                        for (int e = 0; e < 4; ++e) hb[e] = __hmax2(hb[e], zero);  // the ReLU
                        *reinterpret_cast<uint4*>(blk + p_rel[j]) = v[0];
                    }
                    continue;
                }
                for (int item = gtid; item < S4P_PH * S4P_PW * 8; item += 128) {
                    const int q = item & 7, pp = item >> 3;
                    const int pr = pp / S4P_PW, pc = pp - pr * S4P_PW;
                    const int py = ty * S4P_PH + pr, px = tx * S4P_PW + pc;
                    if (py >= p.PH || px >= p.PW) continue;
                    uint4 best = make_uint4(0u, 0u, 0u, 0u);
                    __nv_bfloat162* hb = reinterpret_cast<__nv_bfloat162*>(&best);
                    hb[0] = hb[1] = hb[2] = hb[3] = zero;
#pragma unroll
                    for (int dy = 0; dy < 3; ++dy) {
                        const int lr = 2 * pr + dy;
                        if (r0 + lr < 0 || r0 + lr >= p.OH) continue;
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
                            const int lc = 2 * pc + dx;
                            if (c0 + lc < 0 || c0 + lc >= p.OW) continue;
                            const uint4 v = *reinterpret_cast<const uint4*>(so + swz_off<128>(lr * TW + lc, q));
                            const __nv_bfloat162* hv = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
                            for (int j = 0; j < 4; ++j) hb[j] = __hmax2(hb[j], hv[j]);
                        }
                    }
                    *reinterpret_cast<uint4*>(p.pooled + (((long long)n * p.PH + py) * p.PW + px) * 64 + q * 8) = best;
                }
                continue;  // (this group's next bar_a orders these reads before its staging tile is written again)
            }
            asm volatile("bar.sync 2, 128;" ::: "memory");
            if (leader) {
                const int tx = tile % p.tiles_x;
                const int t2 = tile / p.tiles_x;
                const int ty = t2 % p.tiles_y;
                const int n = t2 / p.tiles_y;
                tma_store_4d(&tmY, so, 0, tx * TW, ty * TH, n);
                bulk_commit();
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the OTHER staging tile is free again
            }
        }
        if (leader) bulk_wait0();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TCOLS>(tmem_base);
    }
}

}  // namespace fod

extern "C" int fod_stem_conv7x7s2_p4(const void* x4p, const void* wk4, const float* bias, void* y, int32_t N, int32_t H,
                                     int32_t W, void* stream) {
    using namespace fod;
    if (!x4p || !wk4 || !bias || !y || N <= 0 || H <= 0 || W <= 0) return fod_fail(FOD_E_ARG, "stem_conv_p4: null/empty");
    if ((W % 8) || (H % 2)) return fod_fail(FOD_E_ARG, "stem_conv_p4: W must be a multiple of 8 and H even");
    if ((reinterpret_cast<uintptr_t>(x4p) | reinterpret_cast<uintptr_t>(wk4) | reinterpret_cast<uintptr_t>(y)) & 15)
        return fod_fail(FOD_E_ARG, "stem_conv_p4: x / wk / y must be 16-byte aligned");
    const int OH = H / 2, OW = W / 2;
    const long long pitch = (long long)(W + 8) * 8;  // bytes per padded image row
    static const bool raw = [] { const char* e = getenv("FOD_STEM_RAW"); return e ? atoi(e) != 0 : true; }();
    const int TW = raw ? 8 : 16, TH = raw ? 16 : 8;
    CUtensorMap tmX, tmY;
    if (raw) {
        uint64_t dims[4] = {(uint64_t)(W + 8) * 4, 2, (uint64_t)H / 2, (uint64_t)N};
        uint64_t str[3] = {(uint64_t)pitch, 2ull * pitch, (uint64_t)H * pitch};
        uint32_t box[4] = {S4R_ROW_BYTES / 2, 1, (uint32_t)S4R_BOX_ROWS, 1};
        if (make_tmap_bf16(&tmX, x4p, 4, dims, str, box, 0)) return fod_fail(FOD_E_TMAP, "stem_conv_p4: tmap X (rows)");
    } else {
        uint64_t dims[5] = {32, (uint64_t)OW, 2, (uint64_t)H / 2, (uint64_t)N};
        uint64_t str[4] = {16, (uint64_t)pitch, 2ull * pitch, (uint64_t)H * pitch};
        uint32_t box[5] = {32, 16, 1, (uint32_t)S4_BOX_ROWS, 1};
        if (make_tmap_bf16(&tmX, x4p, 5, dims, str, box, 64)) return fod_fail(FOD_E_TMAP, "stem_conv_p4: tmap X");
    }
    {
        uint64_t dims[4] = {64, (uint64_t)OW, (uint64_t)OH, (uint64_t)N};
        uint64_t str[3] = {128, (uint64_t)OW * 128, (uint64_t)OH * OW * 128};
        uint32_t box[4] = {64, (uint32_t)TW, (uint32_t)TH, 1};
        if (make_tmap_bf16(&tmY, y, 4, dims, str, box, 128)) return fod_fail(FOD_E_TMAP, "stem_conv_p4: tmap Y");
    }
    Stem4Params p;
    p.wk = reinterpret_cast<const __nv_bfloat16*>(wk4);
    p.bias = bias;
    p.tiles_x = (OW + TW - 1) / TW;
    p.tiles_y = (OH + TH - 1) / TH;
    p.pooled = nullptr;
    p.OH = OH; p.OW = OW; p.PH = p.PW = 0;
    const long long tiles = (long long)N * p.tiles_x * p.tiles_y;
    if (tiles > 0x7fffffffLL) return fod_fail(FOD_E_ARG, "stem_conv_p4: too many tiles");
    p.tiles = (int)tiles;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(stem_conv7x7s2_p4_kernel<true, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, s4_smem<true, false, 1>());
        cudaFuncSetAttribute(stem_conv7x7s2_p4_kernel<false, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, s4_smem<false, false, 1>());
        attr_set = true;
    }
    const int grid = (int)(tiles < 148 ? tiles : 148);
    if (raw) stem_conv7x7s2_p4_kernel<true, false, 1><<<grid, S4_THREADS, s4_smem<true, false, 1>(), reinterpret_cast<cudaStream_t>(stream)>>>(tmX, tmY, p);
    else stem_conv7x7s2_p4_kernel<false, false, 1><<<grid, S4_THREADS, s4_smem<false, false, 1>(), reinterpret_cast<cudaStream_t>(stream)>>>(tmX, tmY, p);
    return fod_check_launch("stem_conv7x7s2_p4_kernel");
}

extern "C" int fod_stem_conv_pool_p4(const void* x4p, const void* wk4, const float* bias, void* y, int32_t N, int32_t H,
                                     int32_t W, void* stream) {
    using namespace fod;
    if (!x4p || !wk4 || !bias || !y || N <= 0 || H <= 0 || W <= 0) return fod_fail(FOD_E_ARG, "stem_conv_pool: null/empty");
    if ((W % 8) || (H % 2)) return fod_fail(FOD_E_ARG, "stem_conv_pool: W must be a multiple of 8 and H even");
    if ((reinterpret_cast<uintptr_t>(x4p) | reinterpret_cast<uintptr_t>(wk4) | reinterpret_cast<uintptr_t>(y)) & 15)
        return fod_fail(FOD_E_ARG, "stem_conv_pool: x / wk / y must be 16-byte aligned");
    const int OH = H / 2, OW = W / 2;
    const long long pitch = (long long)(W + 8) * 8;
    CUtensorMap tmX;
    {
        uint64_t dims[4] = {(uint64_t)(W + 8) * 4, 2, (uint64_t)H / 2, (uint64_t)N};
        uint64_t str[3] = {(uint64_t)pitch, 2ull * pitch, (uint64_t)H * pitch};
        uint32_t box[4] = {S4R_ROW_BYTES / 2, 1, (uint32_t)S4R_BOX_ROWS, 1};
        if (make_tmap_bf16(&tmX, x4p, 4, dims, str, box, 0)) return fod_fail(FOD_E_TMAP, "stem_conv_pool: tmap X");
    }
    Stem4Params p;
    p.wk = reinterpret_cast<const __nv_bfloat16*>(wk4);
    p.bias = bias;
    p.pooled = reinterpret_cast<__nv_bfloat16*>(y);
    p.OH = OH; p.OW = OW;
    p.PH = (OH - 1) / 2 + 1;
    p.PW = (OW - 1) / 2 + 1;
    p.tiles_x = (p.PW + S4P_PW - 1) / S4P_PW;
    p.tiles_y = (p.PH + S4P_PH - 1) / S4P_PH;
    const long long tiles = (long long)N * p.tiles_x * p.tiles_y;
    if (tiles > 0x7fffffffLL) return fod_fail(FOD_E_ARG, "stem_conv_pool: too many tiles");
    p.tiles = (int)tiles;
    static const int groups = [] { const char* e = getenv("FOD_STEM_GROUPS"); return e ? atoi(e) : 3; }();
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(stem_conv7x7s2_p4_kernel<true, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, s4_smem<true, true, 2>());
        cudaFuncSetAttribute(stem_conv7x7s2_p4_kernel<true, true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, s4_smem<true, true, 3>());
        cudaFuncSetAttribute(stem_conv7x7s2_p4_kernel<true, true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, s4_smem<true, true, 4>());
        attr_set = true;
    }
    const int grid = (int)(tiles < 148 ? tiles : 148);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (groups == 2) stem_conv7x7s2_p4_kernel<true, true, 2><<<grid, s4_threads<true, 2>(), s4_smem<true, true, 2>(), st>>>(tmX, tmX, p);
    else if (groups == 4) stem_conv7x7s2_p4_kernel<true, true, 4><<<grid, s4_threads<true, 4>(), s4_smem<true, true, 4>(), st>>>(tmX, tmX, p);
    else stem_conv7x7s2_p4_kernel<true, true, 3><<<grid, s4_threads<true, 3>(), s4_smem<true, true, 3>(), st>>>(tmX, tmX, p);
    return fod_check_launch("stem_conv_pool_p4_kernel");
}

namespace fod {
}  // namespace fod

extern "C" int fod_stem_conv7x7s2(const void* x, const void* wk, const float* bias, void* y, int32_t N, int32_t H, int32_t W,
                                  void* stream) {
    using namespace fod;
    if (!x || !wk || !bias || !y || N <= 0 || H <= 0 || W <= 0) return fod_fail(FOD_E_ARG, "stem_conv: null/empty");
    if (W % 8) return fod_fail(FOD_E_ARG, "stem_conv: W must be a multiple of 8 (16-byte image rows)");
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(wk) | reinterpret_cast<uintptr_t>(y)) & 15)
        return fod_fail(FOD_E_ARG, "stem_conv: x / wk / y must be 16-byte aligned");
    StemParams p;
    p.x = reinterpret_cast<const __nv_bfloat16*>(x);
    p.wk = reinterpret_cast<const __nv_bfloat16*>(wk);
    p.bias = bias;
    p.y = reinterpret_cast<__nv_bfloat16*>(y);
    p.N = N; p.H = H; p.W = W;
    p.OH = (H - 1) / 2 + 1;
    p.OW = (W - 1) / 2 + 1;
    p.tiles_x = (p.OW + ST_TW - 1) / ST_TW;
    p.tiles_y = (p.OH + ST_TH - 1) / ST_TH;
    const long long tiles = (long long)N * p.tiles_x * p.tiles_y;
    if (tiles > 0x7fffffffLL) return fod_fail(FOD_E_ARG, "stem_conv: too many tiles");
    p.tiles = (int)tiles;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(stem_conv7x7s2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM);
        attr_set = true;
    }
    const int grid = (int)(tiles < 148 * 2 ? tiles : 148 * 2);
    stem_conv7x7s2_kernel<<<grid, 128, ST_SMEM, reinterpret_cast<cudaStream_t>(stream)>>>(p);
    return fod_check_launch("stem_conv7x7s2_kernel");
}
