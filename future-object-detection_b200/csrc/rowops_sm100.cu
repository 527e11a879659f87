// HBM-bound row kernels over 256-channel token rows (one warp per row, 16-byte vector loads, warp-shuffle
// reductions; grid sized to 148 SMs x 4 CTAs):
//   fod_add_ln_fwd : y = LayerNorm(x + bcast[row / rows_per_group])   transformer.py:486 (norm_eda), :386 (decoder norm)
//   fod_ln_bwd     : LayerNorm backward (dz, dgamma, dbeta) for every post-norm block, transformer.py:416-419 etc.
//   fod_colsum     : per-group column sums (bias gradients, broadcast-operand gradients)
// Algorithmic bytes per row: fwd 512 B read + 512 B write; bwd 1024 B read + 512 B write (bf16, D = 256).
#include "fod_internal.h"
#include "sm100_ptx.cuh"
#include "dropout.cuh"

namespace fod {

constexpr int RO_D = 256;
constexpr int RO_WARPS = 8;

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]);
    u.y = pack_bf16x2(f[2], f[3]);
    u.z = pack_bf16x2(f[4], f[5]);
    u.w = pack_bf16x2(f[6], f[7]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(RO_WARPS * 32)
add_ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bcast, int rows_per_group,
                  const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                  float* __restrict__ mean_out, float* __restrict__ rstd_out, long long M, float eps,
                  const __nv_bfloat16* x2, __nv_bfloat16* __restrict__ pre_out, const DropoutCfg drop) {
    // x2 (may alias x): a second full-size addend that goes through dropout first — y = LN(x + dropout(x2) [+ bcast]),
    // the "x + dropout(sublayer(x))" of blocks whose sub-layer output is not a GEMM epilogue (transformer.py:117,485);
    // pre_out receives the pre-LayerNorm sum for ln_bwd.  Mask index: row * 32 + lane, as fod_ln_bwd_drop regenerates it.
    pdl_launch_dependents();
    pdl_wait();
    PhiloxKey key = {0u, 0u, 0u, 0u};
    if (drop.state) key = dropout_key(drop);
    const int lane = threadIdx.x & 31;
    const int c0 = lane * 8;
    float g[8], bt[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        g[j] = __ldg(gamma + c0 + j);
        bt[j] = __ldg(beta + c0 + j);
    }
    const long long stride = (long long)gridDim.x * RO_WARPS;
    for (long long row = (long long)blockIdx.x * RO_WARPS + (threadIdx.x >> 5); row < M; row += stride) {
        float z[8];
        load8(x + row * RO_D + c0, z);
        if (x2) {
            float u[8];
            load8(x2 + row * RO_D + c0, u);
            if (drop.state) dropout_apply8(u, key, (unsigned long long)row * (RO_D / 8) + lane, drop.thresh, drop.scale);
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j] += u[j];
        }
        if (bcast) {
            float o[8];
            load8(bcast + (row / rows_per_group) * RO_D + c0, o);
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j] += o[j];
        }
        if (pre_out) store8(pre_out + row * RO_D + c0, z);
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += z[j];
        const float mean = warp_sum(s) * (1.f / RO_D);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float d = z[j] - mean;
            q += d * d;
        }
        const float rstd = rsqrtf(warp_sum(q) * (1.f / RO_D) + eps);
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (z[j] - mean) * rstd * g[j] + bt[j];
        store8(y + row * RO_D + c0, o);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
    }
}

// dz = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma,  xhat = (z - mean) * rstd,
// z = pre (+ bcast[row / rows_per_group]).  dy may carry a second addend dy2 (two consumers of the LN output).
__global__ void __launch_bounds__(RO_WARPS * 32)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ dy2,
              const __nv_bfloat16* __restrict__ pre, const __nv_bfloat16* __restrict__ bcast, int rows_per_group,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const float* __restrict__ gamma,
              __nv_bfloat16* __restrict__ dz, float* __restrict__ dgamma, float* __restrict__ dbeta,
              float* __restrict__ dzsum, long long M, __nv_bfloat16* __restrict__ dz_drop, const DropoutCfg drop) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float red[3][RO_WARPS][RO_D];
    PhiloxKey key = {0u, 0u, 0u, 0u};
    if (drop.state) key = dropout_key(drop);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = lane * 8;
    float g[8], ag[8], ab[8], az[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        g[j] = __ldg(gamma + c0 + j);
        ag[j] = 0.f;
        ab[j] = 0.f;
        az[j] = 0.f;
    }
    const long long stride = (long long)gridDim.x * RO_WARPS;
    for (long long row = (long long)blockIdx.x * RO_WARPS + warp; row < M; row += stride) {
        float d[8], z[8];
        load8(dy + row * RO_D + c0, d);
        if (dy2) {
            float e[8];
            load8(dy2 + row * RO_D + c0, e);
#pragma unroll
            for (int j = 0; j < 8; ++j) d[j] += e[j];
        }
        load8(pre + row * RO_D + c0, z);
        if (bcast) {
            float o[8];
            load8(bcast + (row / rows_per_group) * RO_D + c0, o);
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j] += o[j];
        }
        const float mean = mean_in[row], rstd = rstd_in[row];
        float s1 = 0.f, s2 = 0.f, xh[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            xh[j] = (z[j] - mean) * rstd;
            const float gj = d[j] * g[j];
            s1 += gj;
            s2 += gj * xh[j];
            ag[j] += d[j] * xh[j];
            ab[j] += d[j];
        }
        s1 = warp_sum(s1) * (1.f / RO_D);
        s2 = warp_sum(s2) * (1.f / RO_D);
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (d[j] * g[j] - s1 - xh[j] * s2);
        store8(dz + row * RO_D + c0, o);
        if (drop.state) {  // gradient of the dropped-out addend: this lane's 8 columns are one Philox group (row * 32 + lane)
            dropout_apply8(o, key, (unsigned long long)row * (RO_D / 8) + lane, drop.thresh, drop.scale);
            if (dz_drop) store8(dz_drop + row * RO_D + c0, o);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) az[j] += o[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        red[0][warp][c0 + j] = ag[j];
        red[1][warp][c0 + j] = ab[j];
        red[2][warp][c0 + j] = az[j];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < RO_D; c += RO_WARPS * 32) {
        float sg = 0.f, sb = 0.f, sz = 0.f;
#pragma unroll
        for (int w = 0; w < RO_WARPS; ++w) {
            sg += red[0][w][c];
            sb += red[1][w][c];
            sz += red[2][w][c];
        }
        if (dgamma) atomicAdd(dgamma + c, sg);
        if (dbeta) atomicAdd(dbeta + c, sb);
        if (dzsum) atomicAdd(dzsum + c, sz);  // bias gradient of the linear layer feeding this LayerNorm
    }
}

// out[g][c] += sum over the rows of group g of x[row][c].  A CTA owns a strip of rows of one group and a 256-column
// slab: 32 lanes x 8 columns (one 16-byte load) per row, 8 warps on 8 different rows, 4 independent loads in flight
// per thread; the 8 warps are combined through shared memory and one atomic per column leaves the CTA.
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, float* __restrict__ out, int N, int rows_per_group,
              int strips_per_group, int rows_per_strip, int vec_ok) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float red[8][256];
    const int g = blockIdx.x / strips_per_group;
    const int strip = blockIdx.x % strips_per_group;
    const int r0 = strip * rows_per_strip;
    const int r1 = min(rows_per_group, r0 + rows_per_strip);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = blockIdx.y * 256 + lane * 8;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const __nv_bfloat16* base = x + (long long)g * rows_per_group * ldx + c0;
    if (vec_ok && c0 + 8 <= N) {
        int r = r0 + warp;
        for (; r + 24 < r1; r += 32) {
            float f[4][8];
#pragma unroll
            for (int u = 0; u < 4; ++u) load8(base + (long long)(r + 8 * u) * ldx, f[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] += f[u][j];
        }
        for (; r < r1; r += 8) {
            float f[8];
            load8(base + (long long)r * ldx, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
    } else {
        for (int r = r0 + warp; r < r1; r += 8)
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (c0 + j < N) acc[j] += __bfloat162float(base[(long long)r * ldx + j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    const int c = blockIdx.y * 256 + threadIdx.x;
    if (c < N) {
        float s2 = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s2 += red[w][threadIdx.x];
        atomicAdd(out + (long long)g * N + c, s2);
    }
}

static int ro_grid(long long M) {
    long long need = (M + RO_WARPS - 1) / RO_WARPS;
    long long cap = 148 * 4;
    return (int)(need < cap ? (need < 1 ? 1 : need) : cap);
}

}  // namespace fod

using namespace fod;

extern "C" int fod_add_ln_fwd(const void* x, const void* bcast, int32_t rows_per_group, const float* gamma,
                              const float* beta, void* y, float* mean_out, float* rstd_out, int64_t M, float eps,
                              void* stream) {
    if (!x || !gamma || !beta || !y || M <= 0) return fod_fail(FOD_E_ARG, "add_ln_fwd: null/empty");
    if (bcast && rows_per_group <= 0) return fod_fail(FOD_E_ARG, "add_ln_fwd: rows_per_group");
    fod_launch(add_ln_fwd_kernel, dim3(ro_grid(M)), dim3(RO_WARPS * 32), 0, reinterpret_cast<cudaStream_t>(stream),
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(bcast), rows_per_group,
        gamma, beta, reinterpret_cast<__nv_bfloat16*>(y), mean_out, rstd_out, M, eps,
        static_cast<const __nv_bfloat16*>(nullptr), static_cast<__nv_bfloat16*>(nullptr), DropoutCfg{nullptr, 0u, 0u, 1.f});
    return fod_check_launch("add_ln_fwd_kernel");
}

extern "C" int fod_add_ln_fwd_drop(const void* x, const void* x2, const void* bcast, int32_t rows_per_group,
                                   const float* gamma, const float* beta, void* y, void* pre_out, float* mean_out,
                                   float* rstd_out, int64_t M, float eps, const fod_dropout* drop, void* stream) {
    if (!x || !x2 || !gamma || !beta || !y || M <= 0) return fod_fail(FOD_E_ARG, "add_ln_fwd_drop: null/empty");
    if (bcast && rows_per_group <= 0) return fod_fail(FOD_E_ARG, "add_ln_fwd_drop: rows_per_group");
    DropoutCfg d = {nullptr, 0u, 0u, 1.f};
    if (drop && drop->rng_state && drop->p > 0.f) {
        if (!(drop->p < 1.f)) return fod_fail(FOD_E_ARG, "add_ln_fwd_drop: dropout p must be in [0, 1)");
        d.state = reinterpret_cast<const unsigned long long*>(drop->rng_state);
        d.call = (unsigned int)drop->call;
        d.thresh = (unsigned int)(drop->p * 65536.f + 0.5f);
        d.scale = 1.f / (1.f - drop->p);
    }
    fod_launch(add_ln_fwd_kernel, dim3(ro_grid(M)), dim3(RO_WARPS * 32), 0, reinterpret_cast<cudaStream_t>(stream),
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(bcast), rows_per_group,
        gamma, beta, reinterpret_cast<__nv_bfloat16*>(y), mean_out, rstd_out, M, eps,
        reinterpret_cast<const __nv_bfloat16*>(x2), reinterpret_cast<__nv_bfloat16*>(pre_out), d);
    return fod_check_launch("add_ln_fwd_kernel");
}

extern "C" int fod_ln_bwd_drop(const void* dy, const void* dy2, const void* pre, const void* bcast, int32_t rows_per_group,
                               const float* mean, const float* rstd, const float* gamma, void* dz, void* dz_drop,
                               float* dgamma, float* dbeta, float* dzsum, int64_t M, const fod_dropout* drop, void* stream) {
    if (!dy || !pre || !mean || !rstd || !gamma || !dz || M <= 0) return fod_fail(FOD_E_ARG, "ln_bwd: null/empty");
    if (bcast && rows_per_group <= 0) return fod_fail(FOD_E_ARG, "ln_bwd: rows_per_group");
    DropoutCfg d = {nullptr, 0u, 0u, 1.f};
    if (drop && drop->rng_state && drop->p > 0.f) {
        if (!(drop->p < 1.f)) return fod_fail(FOD_E_ARG, "ln_bwd: dropout p must be in [0, 1)");
        d.state = reinterpret_cast<const unsigned long long*>(drop->rng_state);
        d.call = (unsigned int)drop->call;
        d.thresh = (unsigned int)(drop->p * 65536.f + 0.5f);
        d.scale = 1.f / (1.f - drop->p);
    }
    fod_launch(ln_bwd_kernel, dim3(ro_grid(M)), dim3(RO_WARPS * 32), 0, reinterpret_cast<cudaStream_t>(stream),
        reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(dy2),
        reinterpret_cast<const __nv_bfloat16*>(pre), reinterpret_cast<const __nv_bfloat16*>(bcast), rows_per_group, mean,
        rstd, gamma, reinterpret_cast<__nv_bfloat16*>(dz), dgamma, dbeta, dzsum, M, reinterpret_cast<__nv_bfloat16*>(dz_drop), d);
    return fod_check_launch("ln_bwd_kernel");
}

extern "C" int fod_ln_bwd(const void* dy, const void* dy2, const void* pre, const void* bcast, int32_t rows_per_group,
                          const float* mean, const float* rstd, const float* gamma, void* dz, float* dgamma,
                          float* dbeta, float* dzsum, int64_t M, void* stream) {
    return fod_ln_bwd_drop(dy, dy2, pre, bcast, rows_per_group, mean, rstd, gamma, dz, nullptr, dgamma, dbeta, dzsum, M, nullptr,
                           stream);
}

namespace fod {
__global__ void rng_advance_kernel(unsigned long long* state) { state[1] += 1ull; }
}  // namespace fod

extern "C" int fod_rng_advance(void* rng_state, void* stream) {
    if (!rng_state) return fod_fail(FOD_E_ARG, "rng_advance: null");
    fod::rng_advance_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(rng_state));
    return fod_check_launch("rng_advance_kernel");
}

extern "C" int fod_colsum(const void* x, int64_t ldx, float* out, int64_t M, int32_t N, int32_t rows_per_group,
                          void* stream) {
    if (!x || !out || M <= 0 || N <= 0) return fod_fail(FOD_E_ARG, "colsum: null/empty");
    if (rows_per_group <= 0) rows_per_group = (int32_t)M;
    if (M % rows_per_group) return fod_fail(FOD_E_ARG, "colsum: M must be a multiple of rows_per_group");
    const int groups = (int)(M / rows_per_group);
    const int col_slabs = (N + 255) / 256;
    int strips = (148 * 2 + groups * col_slabs - 1) / (groups * col_slabs);
    const int max_strips = (rows_per_group + 63) / 64;
    if (strips > max_strips) strips = max_strips;
    if (strips < 1) strips = 1;
    const int rows_per_strip = (rows_per_group + strips - 1) / strips;
    strips = (rows_per_group + rows_per_strip - 1) / rows_per_strip;
    dim3 grid(groups * strips, col_slabs);
    const int vec_ok = (ldx % 8 == 0) && !(reinterpret_cast<uintptr_t>(x) & 15);
    fod_launch(colsum_kernel, grid, dim3(256), 0, reinterpret_cast<cudaStream_t>(stream),
        reinterpret_cast<const __nv_bfloat16*>(x), ldx, out, N, rows_per_group, strips, rows_per_strip, vec_ok);
    return fod_check_launch("colsum_kernel");
}

// ---------------------------------------------------------------------------------------------------------------
// 3x3 / stride 2 / pad 1 max pooling over NHWC bf16 (the ResNet stem pool, torchvision resnet.py `maxpool`, reached
// through ConditionalDETR/models/backbone.py:88-96): one thread = 8 channels (16 bytes) of one output pixel, nine
// 16-byte loads.  HBM-bound: reads the input once (window overlap is served by L1/L2), writes a quarter of it.
namespace fod {
__global__ void __launch_bounds__(256)
maxpool3x3s2_nhwc_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int H, int W, int C8,
                         int OH, int OW) {
    const long long total = (long long)N * OH * OW * C8;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int c8 = (int)(i % C8);
        long long t = i / C8;
        const int ow = (int)(t % OW);
        t /= OW;
        const int oh = (int)(t % OH);
        const int n = (int)(t / OH);
        __nv_bfloat162 m[4];
        const __nv_bfloat162 ninf = __floats2bfloat162_rn(-INFINITY, -INFINITY);
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = ninf;
#pragma unroll
        for (int dh = 0; dh < 3; ++dh) {
            const int ih = oh * 2 - 1 + dh;
            if (ih < 0 || ih >= H) continue;
#pragma unroll
            for (int dw = 0; dw < 3; ++dw) {
                const int iw = ow * 2 - 1 + dw;
                if (iw < 0 || iw >= W) continue;
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + (((long long)n * H + ih) * W + iw) * C8 * 8) + c8);
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], h[j]);
            }
        }
        uint4 o;
        __nv_bfloat162* oh2 = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j) oh2[j] = m[j];
        reinterpret_cast<uint4*>(y)[i] = o;
    }
}
}  // namespace fod

extern "C" int fod_maxpool3x3s2_nhwc(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, void* stream) {
    if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0 || C % 8) return fod_fail(FOD_E_ARG, "maxpool: null/empty or C % 8 != 0");
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) return fod_fail(FOD_E_ARG, "maxpool: alignment");
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    const long long total = (long long)N * OH * OW * (C / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    fod::maxpool3x3s2_nhwc_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), N, H, W, C / 8, OH, OW);
    return fod::fod_check_launch("maxpool3x3s2_nhwc_kernel");
}

// ---------------------------------------------------------------- sum over the leading (batch) dimension
// out[i] = sum_g x[g * n + i]: the gradient of a batch-invariant operand that entered a GEMM as a row-modular residual
// (positional projections shared by every sequence, transformer.py:459-460,160-166) or an attention call as a broadcast
// part (query_sine / key_pos).  Thread = 8 consecutive elements, fp32 accumulation over g with 8 loads in flight.
namespace fod {
__global__ void __launch_bounds__(256)
batch_sum_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, long long G, long long n8) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        const uint4* src = reinterpret_cast<const uint4*>(x) + i;
        long long g = 0;
        for (; g + 8 <= G; g += 8) {
            uint4 u[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) u[k] = __ldg(src + (g + k) * n8);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u[k]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 t = __bfloat1622float2(h[j]);
                    acc[2 * j] += t.x;
                    acc[2 * j + 1] += t.y;
                }
            }
        }
        for (; g < G; ++g) {
            const uint4 u = __ldg(src + g * n8);
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(h[j]);
                acc[2 * j] += t.x;
                acc[2 * j + 1] += t.y;
            }
        }
        store8(out + i * 8, acc);
    }
}
}  // namespace fod

extern "C" int fod_batch_sum(const void* x, void* out, int64_t G, int64_t n, void* stream) {
    if (!x || !out || G <= 0 || n <= 0 || n % 8) return fod_fail(FOD_E_ARG, "batch_sum: null/empty or n % 8 != 0");
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) return fod_fail(FOD_E_ARG, "batch_sum: alignment");
    const long long n8 = n / 8;
    long long blocks = (n8 + 255) / 256;
    if (blocks > 148LL * 8) blocks = 148LL * 8;
    fod::batch_sum_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), G, n8);
    return fod::fod_check_launch("batch_sum_kernel");
}

// ---------------------------------------------------------------- input frames: fp32 planar -> bf16 NHWC
// The model receives `video` (B, L, 3, H, W) fp32 and drops the last frame (paper.py:463-468); the backbone wants the
// remaining B*(L-1) frames as channels-last bf16.  torch does this as three full passes (frame-slice copy, channels-last
// copy, cast: 188 + 199 + 59 us at batch 32 and 448x800); here it is one: a thread takes 8 consecutive pixels of a row,
// reads 2 x 16 bytes from each of the 3 planes and writes 48 contiguous bytes (pixel-major).  HBM-bound: 18 B per pixel.
namespace fod {
__global__ void __launch_bounds__(256)
frames_to_nhwc_bf16_kernel(const float* __restrict__ video, long long stride_b, long long stride_t, int T, int H, int W,
                           __nv_bfloat16* __restrict__ out, long long groups) {
    const long long hw = (long long)H * W;
    const int wg = W / 8;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (long long)gridDim.x * blockDim.x) {
        const int xg = (int)(g % wg);
        const long long r = g / wg;          // frame * H + row
        const int y = (int)(r % H);
        const long long n = r / H;           // frame index b * T + t
        const float* src = video + (n / T) * stride_b + (n % T) * stride_t + (long long)y * W + xg * 8;
        float c[3][8];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(src + ch * hw));
            const float4 b = __ldg(reinterpret_cast<const float4*>(src + ch * hw) + 1);
            c[ch][0] = a.x; c[ch][1] = a.y; c[ch][2] = a.z; c[ch][3] = a.w;
            c[ch][4] = b.x; c[ch][5] = b.y; c[ch][6] = b.z; c[ch][7] = b.w;
        }
        float f[24];
#pragma unroll
        for (int px = 0; px < 8; ++px)
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) f[px * 3 + ch] = c[ch][px];
        uint4* dst = reinterpret_cast<uint4*>(out + ((n * H + y) * (long long)W + xg * 8) * 3);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            uint4 u;
            u.x = pack_bf16x2(f[q * 8 + 0], f[q * 8 + 1]);
            u.y = pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
            u.z = pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]);
            u.w = pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
            dst[q] = u;
        }
    }
}
}  // namespace fod

extern "C" int fod_frames_to_nhwc_bf16(const float* video, int64_t stride_b, int64_t stride_t, int32_t B, int32_t T,
                                       int32_t H, int32_t W, void* out, void* stream) {
    if (!video || !out || B <= 0 || T <= 0 || H <= 0 || W <= 0) return fod_fail(FOD_E_ARG, "frames_to_nhwc: null/empty");
    if (W % 8 || stride_b % 4 || stride_t % 4 || ((long long)H * W) % 4 ||
        ((reinterpret_cast<uintptr_t>(video) | reinterpret_cast<uintptr_t>(out)) & 15))
        return fod_fail(FOD_E_ARG, "frames_to_nhwc: W % 8 == 0 and 16-byte aligned planes / frames needed");
    const long long groups = (long long)B * T * H * (W / 8);
    long long blocks = (groups + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    fod::frames_to_nhwc_bf16_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        video, stride_b, stride_t, T, H, W, reinterpret_cast<__nv_bfloat16*>(out), groups);
    return fod::fod_check_launch("frames_to_nhwc_bf16_kernel");
}

// The same frames as (N, H, W + 8, 4) bf16: pixel x at padded column x + 3, channel 3 and the 3 + 5 border columns zero —
// the layout fod_stem_conv7x7s2_p4 reads through one overlapping-window tensor map (csrc/stemconv_sm100.cu).  One
// thread writes 8 padded columns (64 contiguous bytes); every element of `out` is written, borders included.
namespace fod {
template <bool VEC>
__global__ void __launch_bounds__(256)
frames_to_nhwc4p_bf16_kernel(const float* __restrict__ video, long long stride_b, long long stride_t, int T, int H, int W,
                             __nv_bfloat16* __restrict__ out, long long groups) {
    const long long hw = (long long)H * W;
    const int wg = (W + 8) / 8;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (long long)gridDim.x * blockDim.x) {
        const int xg = (int)(g % wg);
        const long long r = g / wg;          // frame * H + row
        const int y = (int)(r % H);
        const long long n = r / H;           // frame index b * T + t
        const float* src = video + (n / T) * stride_b + (n % T) * stride_t + (long long)y * W;
        const int x0 = xg * 8 - 3;
        uint4* dst = reinterpret_cast<uint4*>(out + (r * (long long)(W + 8) + xg * 8) * 4);
        float c[3][8];  // image columns x0 .. x0 + 7 of the three planes
        if (VEC) {      // aligned 16-byte loads of columns [8 xg - 4, 8 xg + 8): the wanted eight are elements 1..8
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float4* pl = reinterpret_cast<const float4*>(src + ch * hw + xg * 8);
                const float4 a = xg > 0 ? __ldg(pl - 1) : z4;
                const float4 b = xg * 8 < W ? __ldg(pl) : z4;
                const float4 d = xg * 8 < W ? __ldg(pl + 1) : z4;
                c[ch][0] = a.y; c[ch][1] = a.z; c[ch][2] = a.w;
                c[ch][3] = b.x; c[ch][4] = b.y; c[ch][5] = b.z; c[ch][6] = b.w;
                c[ch][7] = d.x;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int x = x0 + i;
                const bool ok = x >= 0 && x < W;
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) c[ch][i] = ok ? __ldg(src + ch * hw + x) : 0.f;
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {  // two pixels per 16-byte store
            uint4 u;
            u.x = pack_bf16x2(c[0][2 * q], c[1][2 * q]);
            u.y = pack_bf16x2(c[2][2 * q], 0.f);
            u.z = pack_bf16x2(c[0][2 * q + 1], c[1][2 * q + 1]);
            u.w = pack_bf16x2(c[2][2 * q + 1], 0.f);
            dst[q] = u;
        }
    }
}
}  // namespace fod

extern "C" int fod_frames_to_nhwc4p_bf16(const float* video, int64_t stride_b, int64_t stride_t, int32_t B, int32_t T,
                                         int32_t H, int32_t W, void* out, void* stream) {
    if (!video || !out || B <= 0 || T <= 0 || H <= 0 || W <= 0) return fod_fail(FOD_E_ARG, "frames_to_nhwc4p: null/empty");
    if (W % 8 || (reinterpret_cast<uintptr_t>(out) & 15)) return fod_fail(FOD_E_ARG, "frames_to_nhwc4p: W % 8 == 0 and a 16-byte aligned output needed");
    const long long groups = (long long)B * T * H * ((W + 8) / 8);
    long long blocks = (groups + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    const bool vec = !(stride_b % 4) && !(stride_t % 4) && !(((long long)H * W) % 4) && !(reinterpret_cast<uintptr_t>(video) & 15);
    if (vec)
        fod::frames_to_nhwc4p_bf16_kernel<true><<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
            video, stride_b, stride_t, T, H, W, reinterpret_cast<__nv_bfloat16*>(out), groups);
    else
        fod::frames_to_nhwc4p_bf16_kernel<false><<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
            video, stride_b, stride_t, T, H, W, reinterpret_cast<__nv_bfloat16*>(out), groups);
    return fod::fod_check_launch("frames_to_nhwc4p_bf16_kernel");
}

// ---------------------------------------------------------------- fused ReLU backward (+ gradient add) for the backbone
// out = (a [+ b]) * (y > 0), flat bf16 arrays of n elements (n % 8 == 0, identical layouts).  In a bottleneck's backward
// the gradient of the block output arrives as two addends (the conv1 input gradient and the identity-path gradient of
// the NEXT block, torchvision resnet.py Bottleneck.forward / ConditionalDETR/models/backbone.py:88-96): autograd would add
// them in one pass (3 tensors of traffic) and mask them in another (3 more); this does both in one (4).  HBM-bound:
// 8 bytes per element with b, 6 without.
namespace fod {
__global__ void __launch_bounds__(256)
relu_bwd_add_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint4* __restrict__ y, uint4* __restrict__ out,
                    long long n8) {
    const __nv_bfloat162 zero = __floats2bfloat162_rn(0.f, 0.f);
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n8; i += (long long)gridDim.x * 256) {
        uint4 ua = __ldg(a + i);
        const uint4 uy = __ldg(y + i);
        __nv_bfloat162* ha = reinterpret_cast<__nv_bfloat162*>(&ua);
        const __nv_bfloat162* hy = reinterpret_cast<const __nv_bfloat162*>(&uy);
        if (b) {
            const uint4 ub = __ldg(b + i);
            const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&ub);
#pragma unroll
            for (int j = 0; j < 4; ++j) ha[j] = __hadd2(ha[j], hb[j]);  // bf16 add, as torch's accumulation of two bf16 gradients
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 m = __hgt2(hy[j], zero);  // 1.0 where y > 0 (a NaN activation masks to 0, as threshold_backward)
            ha[j] = __hmul2(ha[j], m);
        }
        out[i] = ua;
    }
}
}  // namespace fod

extern "C" int fod_relu_bwd_add(const void* a, const void* b, const void* y, void* out, int64_t n, void* stream) {
    if (!a || !y || !out || n <= 0 || n % 8) return fod::fod_fail(FOD_E_ARG, "relu_bwd_add: null/empty or n % 8 != 0");
    if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(out)) & 15)
        return fod::fod_fail(FOD_E_ARG, "relu_bwd_add: 16-byte alignment");
    const long long n8 = n / 8;
    long long blocks = (n8 + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    fod::relu_bwd_add_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), reinterpret_cast<const uint4*>(y),
        reinterpret_cast<uint4*>(out), n8);
    return fod::fod_check_launch("relu_bwd_add_kernel");
}
