// Frozen-BatchNorm weight folding for the cuDNN backbone (the convolutions themselves stay library
// code).  ConditionalDETR/models/backbone.py:53-63 applies y = conv(x, W)·s + t with frozen (s, t); folded, the conv
// runs on W·s and the backward needs dW = dW_folded·s.  Per trainable conv and step that was a chain of small torch
// launches (fp32 multiply, bf16 cast, channels_last copy; fp32 cast + addcmul): ~200 launches per step.  Here:
//   fold_conv_weights_kernel : ONE launch for every trainable conv: out_bf16[o][hw][i] = bf16(w[o][i][hw] * s[o])
//   conv_wgrad_unfold_kernel : one launch per conv: grad[o][i][hw] += float(dwq[o][hw][i]) * s[o]
// Both are HBM-bound streams over the backbone weights (23 M parameters: ~140 MB per fold pass).
#include "fod_internal.h"
#include "sm100_ptx.cuh"

namespace fod {

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

struct FoldDesc {          // mirrors fod_fold_desc
    const float* w;        // (O, I, KH, KW) fp32 contiguous
    const float* scale;    // (O)
    __nv_bfloat16* out;    // (O, KH, KW, I) bf16 = torch channels_last storage of an (O, I, KH, KW) tensor
    int O, I, HW, pad;
};

__global__ void __launch_bounds__(256)
fold_conv_weights_kernel(const FoldDesc* __restrict__ descs) {
    const FoldDesc d = descs[blockIdx.y];
    const long long per_o = (long long)d.I * d.HW;
    const long long n = (long long)d.O * per_o;
    if (d.HW == 1) {  // 1x1: both layouts coincide; 4 elements per thread
        const long long n4 = n >> 2;  // I % 4 == 0 for every ResNet 1x1 conv (checked on the host)
        for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(d.w) + i);
            const float s = __ldg(d.scale + (i << 2) / per_o);
            uint2 u;
            u.x = pack_bf16x2(v.x * s, v.y * s);
            u.y = pack_bf16x2(v.z * s, v.w * s);
            reinterpret_cast<uint2*>(d.out)[i] = u;
        }
    } else {  // KxK: out index (o, hw, i) <- in index (o, i, hw); writes coalesced, reads served by L1/L2
        for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
            const int o = (int)(e / per_o);
            const int r = (int)(e - (long long)o * per_o);
            const int hw = r / d.I, i = r - hw * d.I;
            d.out[e] = __float2bfloat16(__ldg(d.w + (long long)o * per_o + (long long)i * d.HW + hw) * __ldg(d.scale + o));
        }
    }
}

__global__ void __launch_bounds__(256)
conv_wgrad_unfold_kernel(const __nv_bfloat16* __restrict__ dwq, const float* __restrict__ scale, float* __restrict__ grad,
                         int O, int I, int HW) {
    const long long per_o = (long long)I * HW;
    const long long n = (long long)O * per_o;
    if (HW == 1) {
        const long long n4 = n >> 2;
        for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
            const uint2 u = __ldg(reinterpret_cast<const uint2*>(dwq) + i);
            const float s = __ldg(scale + (i << 2) / per_o);
            float4 g = reinterpret_cast<float4*>(grad)[i];
            g.x += bf16_lo(u.x) * s; g.y += bf16_hi(u.x) * s; g.z += bf16_lo(u.y) * s; g.w += bf16_hi(u.y) * s;
            reinterpret_cast<float4*>(grad)[i] = g;
        }
    } else {  // grad index (o, i, hw) <- dwq index (o, hw, i)
        for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
            const int o = (int)(e / per_o);
            const int r = (int)(e - (long long)o * per_o);
            const int i = r / HW, hw = r - i * HW;
            grad[e] += __bfloat162float(dwq[(long long)o * per_o + (long long)hw * I + i]) * __ldg(scale + o);
        }
    }
}

}  // namespace fod

extern "C" int fod_fold_conv_weights(const fod_fold_desc* descs_device, int32_t count, int64_t max_elems, void* stream) {
    if (!descs_device || count <= 0 || max_elems <= 0) return fod::fod_fail(FOD_E_ARG, "fold_conv_weights: null/empty");
    static_assert(sizeof(fod::FoldDesc) == sizeof(fod_fold_desc), "fod_fold_desc layout");
    long long bx = (max_elems / 4 + 255) / 256;
    if (bx > 148) bx = 148;  // grid (bx, count): small convs finish early, the large ones grid-stride
    if (bx < 1) bx = 1;
    dim3 grid((unsigned)bx, (unsigned)count);
    fod::fold_conv_weights_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const fod::FoldDesc*>(descs_device));
    return fod::fod_check_launch("fold_conv_weights_kernel");
}

extern "C" int fod_conv_wgrad_unfold(const void* dwq, const float* scale, float* grad, int32_t O, int32_t I, int32_t HW,
                                     void* stream) {
    if (!dwq || !scale || !grad || O <= 0 || I <= 0 || HW <= 0) return fod::fod_fail(FOD_E_ARG, "conv_wgrad_unfold: null/empty");
    if (HW == 1 && (I % 4 || ((reinterpret_cast<uintptr_t>(dwq) & 7) | (reinterpret_cast<uintptr_t>(grad) & 15))))
        return fod::fod_fail(FOD_E_ARG, "conv_wgrad_unfold: 1x1 path needs I % 4 == 0 and 8/16-byte aligned dwq/grad");
    const long long n = (long long)O * I * HW;
    long long blocks = ((HW == 1 ? n / 4 : n) + 255) / 256;
    if (blocks > 148LL * 8) blocks = 148LL * 8;
    fod::conv_wgrad_unfold_kernel<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(dwq), scale, grad, O, I, HW);
    return fod::fod_check_launch("conv_wgrad_unfold_kernel");
}
