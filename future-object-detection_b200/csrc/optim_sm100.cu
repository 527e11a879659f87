// Flat-buffer optimizer step: gradient-norm clipping + AdamW + bf16 weight shadow in two launches
// over ONE contiguous fp32 parameter / gradient / moment arena, replacing torch.nn.utils.clip_grad_norm_ (trainer.py:186),
// torch.optim.AdamW.step (trainer.py:188; runs/_helper.py:84-107) and the per-weight fp32->bf16 casts of the kernels'
// operands (hundreds of small launches per step).  HBM-bound: sumsq reads 4 B/element; the update reads 16 B
// (p, g, m, v) and writes 14-18 B (p, m, v, bf16 shadow, optional g = 0) per element.
#include "fod_internal.h"
#include "sm100_ptx.cuh"

namespace fod {

constexpr int OPT_THREADS = 256;
constexpr int OPT_MAX_BLOCKS = 148 * 8;

// stage 1: per-block partial sums of squares (deterministic two-stage reduction, no atomics)
__global__ void __launch_bounds__(OPT_THREADS)
sumsq_partial_kernel(const float* __restrict__ g, long long n, float* __restrict__ partial) {
    float acc = 0.f;
    const long long n4 = n >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long long i = (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n4; i += (long long)gridDim.x * OPT_THREADS) {
        const float4 v = __ldg(g4 + i);
        acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const float v = g[(n4 << 2) + threadIdx.x];
        acc += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float red[OPT_THREADS / 32];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < OPT_THREADS / 32; ++w) s += red[w];
        partial[blockIdx.x] = s;
    }
}

// stage 2 (one block): total norm, clip coefficient, step counter and bias corrections -> state[0..7]
//   state[0] = total grad norm, [1] = clip coefficient, [2] = step t (as float), [3] = 1 - beta1^t, [4] = 1 - beta2^t
__global__ void __launch_bounds__(OPT_THREADS)
optim_prepare_kernel(const float* __restrict__ partial, int nblocks, float* __restrict__ state, float max_norm,
                     float beta1, float beta2) {
    float acc = 0.f;
    for (int i = threadIdx.x; i < nblocks; i += OPT_THREADS) acc += partial[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float red[OPT_THREADS / 32];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < OPT_THREADS / 32; ++w) s += red[w];
        const float norm = sqrtf(s);
        const float t = state[2] + 1.f;
        state[0] = norm;
        state[1] = max_norm > 0.f ? fminf(1.f, max_norm / (norm + 1e-6f)) : 1.f;  // torch clip_grad_norm_: clamp(max/(norm+1e-6), max=1)
        state[2] = t;
        state[3] = 1.f - powf(beta1, t);
        state[4] = 1.f - powf(beta2, t);
    }
}

struct AdamSeg {  // parameter group: elements [begin, end) share lr / weight decay
    long long begin, end;
    float lr, wd;
};
struct AdamArgs {
    float* p;
    float* g;
    float* m;
    float* v;
    __nv_bfloat16* shadow;
    const float* state;
    long long n;
    AdamSeg seg[4];
    int nseg;
    int zero_grad;
    float beta1, beta2, eps;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr, float wd, float b1, float b2,
                                         float eps, float step_size_over_lr, float inv_sqrt_bc2) {
    // torch.optim.AdamW (fused, non-amsgrad): decoupled decay, then the Adam update
    p = p * (1.f - lr * wd);
    m = m + (g - m) * (1.f - b1);            // exp_avg.lerp_(grad, 1 - beta1)
    v = b2 * v + (1.f - b2) * g * g;
    const float denom = sqrtf(v) * inv_sqrt_bc2 + eps;
    p = p - (lr * step_size_over_lr) * (m / denom);
}

__global__ void __launch_bounds__(OPT_THREADS)
adamw_flat_kernel(const AdamArgs a) {
    const float clip = a.state[1];
    const float bc1 = a.state[3], bc2 = a.state[4];
    const float inv_bc1 = 1.f / bc1;
    const float inv_sqrt_bc2 = 1.f / sqrtf(bc2);
    const long long n4 = a.n >> 2;  // segments and n are multiples of 4 elements (16-byte slots)
    for (long long i = (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n4; i += (long long)gridDim.x * OPT_THREADS) {
        const long long e = i << 2;
        float lr = a.seg[0].lr, wd = a.seg[0].wd;
#pragma unroll
        for (int s = 1; s < 4; ++s)
            if (s < a.nseg && e >= a.seg[s].begin) {
                lr = a.seg[s].lr;
                wd = a.seg[s].wd;
            }
        float4 p = reinterpret_cast<float4*>(a.p)[i];
        float4 g = reinterpret_cast<const float4*>(a.g)[i];
        float4 m = reinterpret_cast<float4*>(a.m)[i];
        float4 v = reinterpret_cast<float4*>(a.v)[i];
        g.x *= clip; g.y *= clip; g.z *= clip; g.w *= clip;
        adam_one(p.x, g.x, m.x, v.x, lr, wd, a.beta1, a.beta2, a.eps, inv_bc1, inv_sqrt_bc2);
        adam_one(p.y, g.y, m.y, v.y, lr, wd, a.beta1, a.beta2, a.eps, inv_bc1, inv_sqrt_bc2);
        adam_one(p.z, g.z, m.z, v.z, lr, wd, a.beta1, a.beta2, a.eps, inv_bc1, inv_sqrt_bc2);
        adam_one(p.w, g.w, m.w, v.w, lr, wd, a.beta1, a.beta2, a.eps, inv_bc1, inv_sqrt_bc2);
        reinterpret_cast<float4*>(a.p)[i] = p;
        reinterpret_cast<float4*>(a.m)[i] = m;
        reinterpret_cast<float4*>(a.v)[i] = v;
        if (a.shadow) {
            uint2 u;
            u.x = pack_bf16x2(p.x, p.y);
            u.y = pack_bf16x2(p.z, p.w);
            reinterpret_cast<uint2*>(a.shadow)[i] = u;
        }
        if (a.zero_grad) reinterpret_cast<float4*>(a.g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        else reinterpret_cast<float4*>(a.g)[i] = g;  // leave the clipped gradient, as clip_grad_norm_ does
    }
}

__global__ void __launch_bounds__(OPT_THREADS)
cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n4) {
    for (long long i = (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n4; i += (long long)gridDim.x * OPT_THREADS) {
        const float4 p = __ldg(reinterpret_cast<const float4*>(src) + i);
        uint2 u;
        u.x = pack_bf16x2(p.x, p.y);
        u.y = pack_bf16x2(p.z, p.w);
        reinterpret_cast<uint2*>(dst)[i] = u;
    }
}

static int opt_grid(long long n4) {
    long long need = (n4 + OPT_THREADS - 1) / OPT_THREADS;
    return (int)(need < 1 ? 1 : (need > OPT_MAX_BLOCKS ? OPT_MAX_BLOCKS : need));
}

}  // namespace fod

using namespace fod;

extern "C" int fod_optim_workspace_floats(void) { return OPT_MAX_BLOCKS; }

extern "C" int fod_flat_adamw_step(const fod_adamw_args* a, void* stream) {
    if (!a || !a->params || !a->grads || !a->exp_avg || !a->exp_avg_sq || !a->state || !a->partial || a->n <= 0)
        return fod_fail(FOD_E_ARG, "adamw: null/empty");
    if (a->n % 4 || a->num_groups < 1 || a->num_groups > 4) return fod_fail(FOD_E_ARG, "adamw: n % 4 != 0 or 1..4 groups");
    if ((reinterpret_cast<uintptr_t>(a->params) | reinterpret_cast<uintptr_t>(a->grads) |
         reinterpret_cast<uintptr_t>(a->exp_avg) | reinterpret_cast<uintptr_t>(a->exp_avg_sq)) & 15)
        return fod_fail(FOD_E_ARG, "adamw: buffers must be 16-byte aligned");
    if (a->shadow_bf16 && (reinterpret_cast<uintptr_t>(a->shadow_bf16) & 7)) return fod_fail(FOD_E_ARG, "adamw: shadow alignment");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int gb = opt_grid(a->n >> 2);
    sumsq_partial_kernel<<<gb, OPT_THREADS, 0, st>>>(a->grads, a->n, a->partial);
    optim_prepare_kernel<<<1, OPT_THREADS, 0, st>>>(a->partial, gb, a->state, a->max_norm, a->beta1, a->beta2);
    AdamArgs k;
    k.p = a->params; k.g = a->grads; k.m = a->exp_avg; k.v = a->exp_avg_sq;
    k.shadow = reinterpret_cast<__nv_bfloat16*>(a->shadow_bf16);
    k.state = a->state; k.n = a->n; k.nseg = a->num_groups; k.zero_grad = a->zero_grad;
    k.beta1 = a->beta1; k.beta2 = a->beta2; k.eps = a->eps;
    for (int s = 0; s < 4; ++s) {
        k.seg[s].begin = s < a->num_groups ? a->group_begin[s] : a->n;
        k.seg[s].end = a->n;
        k.seg[s].lr = s < a->num_groups ? a->group_lr[s] : 0.f;
        k.seg[s].wd = s < a->num_groups ? a->group_wd[s] : 0.f;
        if (s < a->num_groups && (a->group_begin[s] % 4 || (s && a->group_begin[s] < a->group_begin[s - 1])))
            return fod_fail(FOD_E_ARG, "adamw: group boundaries must be ascending multiples of 4");
    }
    if (a->group_begin[0] != 0) return fod_fail(FOD_E_ARG, "adamw: first group must start at 0");
    adamw_flat_kernel<<<gb, OPT_THREADS, 0, st>>>(k);
    return fod_check_launch("adamw_flat_kernel");
}

extern "C" int fod_cast_bf16(const float* src, void* dst, int64_t n, void* stream) {
    if (!src || !dst || n <= 0 || n % 4) return fod_fail(FOD_E_ARG, "cast_bf16: null/empty or n % 4 != 0");
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 7)) return fod_fail(FOD_E_ARG, "cast_bf16: alignment");
    cast_bf16_kernel<<<opt_grid(n >> 2), OPT_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        src, reinterpret_cast<__nv_bfloat16*>(dst), n >> 2);
    return fod_check_launch("cast_bf16_kernel");
}
