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
#include "fod_internal.h"
#include "sm100_ptx.cuh"

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
