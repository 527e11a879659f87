// Counter-based dropout masks (Philox4x32, Salmon et al. SC'11, with 7 rounds: the smallest round count the paper
// reports as passing BigCrush; the 10 of cuRAND/torch is a safety margin that costs 40% more integer work here) for the train-mode dropout of the reference
// (p = 0.1 on the attention weights, ConditionalDETR/models/attention.py:370 and torch.nn.MultiheadAttention; after every
// sub-block and inside the FFNs, future_od/models/transformer.py:201,207,234,271,285,310,405-412).
//
// No activation-sized mask is stored: forward and backward evaluate the same function (the attention kernels may read
// it from a 1-bit-per-weight cache, fod_attn_keep_bits) of
//   key     = the run's 64-bit seed                      (device memory: fod_rng_state.seed)
//   counter = (element group, step, call site)          step lives in device memory too and is advanced once per
//                                                        training step (fod_rng_advance), so a captured CUDA graph draws
//                                                        fresh masks on every replay; the call site is a host constant
// One Philox call yields 128 bits = eight 16-bit lanes: element e of a tensor uses lane e % 8 of group e / 8 and is KEPT
// when its lane >= round(p * 65536).  RNG streams cannot match torch's, so parity with the reference is statistical
// (keep rate, scaling, variance) and exact at p = 0 — SURVEY.md §7.3.6.
#pragma once
#include <stdint.h>

namespace fod {

struct DropoutCfg {
    const unsigned long long* state;  // device: {seed, step}; nullptr = dropout off
    unsigned int call;                // call-site id (unique per dropout site and layer within a step)
    unsigned int thresh;              // round(p * 65536); an element is dropped when its 16-bit lane < thresh
    float scale;                      // 1 / (1 - p)
};

struct PhiloxKey {
    uint32_t k0, k1, step, call;
};

__device__ __forceinline__ PhiloxKey dropout_key(const DropoutCfg& d) {
    const unsigned long long seed = d.state[0], step = d.state[1];
    PhiloxKey k;
    k.k0 = (uint32_t)seed;
    k.k1 = (uint32_t)(seed >> 32);
    k.step = (uint32_t)step;
    k.call = d.call ^ ((uint32_t)(step >> 32) * 0x9E3779B9u);
    return k;
}

constexpr int kPhiloxRounds = 7;

// 128 random bits for element group `g` (64-bit)
__device__ __forceinline__ uint4 philox4x32(const PhiloxKey& key, unsigned long long g) {
    uint32_t c0 = (uint32_t)g, c1 = (uint32_t)(g >> 32), c2 = key.step, c3 = key.call;
    uint32_t k0 = key.k0, k1 = key.k1;
#pragma unroll
    for (int r = 0; r < kPhiloxRounds; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// keep-mask bits of the 8 elements of group g: bit j set = element 8g + j is kept
__device__ __forceinline__ uint32_t dropout_keep8(const PhiloxKey& key, unsigned long long g, uint32_t thresh) {
    const uint4 r = philox4x32(key, g);
    uint32_t m = 0;
    m |= ((r.x & 0xFFFFu) >= thresh) ? 1u : 0u;
    m |= ((r.x >> 16) >= thresh) ? 2u : 0u;
    m |= ((r.y & 0xFFFFu) >= thresh) ? 4u : 0u;
    m |= ((r.y >> 16) >= thresh) ? 8u : 0u;
    m |= ((r.z & 0xFFFFu) >= thresh) ? 16u : 0u;
    m |= ((r.z >> 16) >= thresh) ? 32u : 0u;
    m |= ((r.w & 0xFFFFu) >= thresh) ? 64u : 0u;
    m |= ((r.w >> 16) >= thresh) ? 128u : 0u;
    return m;
}

// The same decisions without materialising the bit mask: lane j of a group is the low (j even) or high (j odd) half of
// word j / 2, and  half >= thresh  is one unsigned compare of the word (shifted up for the low half) with thresh << 16.
struct Keep8 {
    uint32_t w[4];
    uint32_t th;
    __device__ __forceinline__ bool operator[](int j) const { return ((j & 1) ? w[j >> 1] : (w[j >> 1] << 16)) >= th; }
};
__device__ __forceinline__ Keep8 dropout_draw8(const PhiloxKey& key, unsigned long long g, uint32_t thresh) {
    const uint4 r = philox4x32(key, g);
    Keep8 k;
    k.w[0] = r.x;
    k.w[1] = r.y;
    k.w[2] = r.z;
    k.w[3] = r.w;
    k.th = thresh << 16;
    return k;
}
// z[j] = kept ? z[j] * scale : 0 for the 8 elements of group g
__device__ __forceinline__ void dropout_apply8(float* z, const PhiloxKey& key, unsigned long long g, uint32_t thresh, float scale) {
    const Keep8 k = dropout_draw8(key, g, thresh);
#pragma unroll
    for (int j = 0; j < 8; ++j) z[j] = k[j] ? z[j] * scale : 0.f;
}

}  // namespace fod
