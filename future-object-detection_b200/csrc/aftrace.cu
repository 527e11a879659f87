// Debug tool: timeline of the fused attention backward (CTA 0) from clock64 stamps (library built with -DFOD_AF_TRACE=1).
// Build: make trace ; run on a B200: ./aftrace_fod [B H Lq Lk]
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../../include/fod_b200.h"
extern "C" int fod_debug_af_trace(long long* host_out);

int main(int argc, char** argv) {
    const int B = argc > 1 ? atoi(argv[1]) : 16, H = argc > 2 ? atoi(argv[2]) : 8, Lq = argc > 3 ? atoi(argv[3]) : 350,
              Lk = argc > 4 ? atoi(argv[4]) : 350, C = H * 32;
    std::vector<__nv_bfloat16> hq((size_t)B * Lq * C), hk((size_t)B * Lk * C);
    unsigned s = 1;
    auto rnd = [&] { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 32768.f - 1.f; };
    for (auto& x : hq) x = __float2bfloat16(rnd());
    for (auto& x : hk) x = __float2bfloat16(rnd());
    __nv_bfloat16 *q, *k, *v, *o, *dout, *dq, *dk, *dv;
    float *lse, *delta;
    cudaMalloc(&q, hq.size() * 2); cudaMalloc(&k, hk.size() * 2); cudaMalloc(&v, hk.size() * 2);
    cudaMalloc(&o, hq.size() * 2); cudaMalloc(&dout, hq.size() * 2); cudaMalloc(&dq, hq.size() * 2);
    cudaMalloc(&dk, hk.size() * 2); cudaMalloc(&dv, hk.size() * 2);
    cudaMalloc(&lse, (size_t)B * H * Lq * 4); cudaMalloc(&delta, (size_t)B * H * Lq * 4);
    cudaMemcpy(q, hq.data(), hq.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dout, hq.data(), hq.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(k, hk.data(), hk.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(v, hk.data(), hk.size() * 2, cudaMemcpyHostToDevice);
    fod_attn_bwd_args a = {};
    a.fwd.q1 = q; a.fwd.k1 = k; a.fwd.v = v; a.fwd.out = o; a.fwd.lse = lse;
    a.fwd.ldq1 = a.fwd.ldk1 = a.fwd.ldv = a.fwd.ldo = C;
    a.fwd.bsq1 = a.fwd.bso = (int64_t)Lq * C; a.fwd.bsk1 = a.fwd.bsv = (int64_t)Lk * C;
    a.fwd.B = B; a.fwd.H = H; a.fwd.Lq = Lq; a.fwd.Lk = Lk; a.fwd.scale = 0.17677669f;
    a.dout = dout; a.lddo = C; a.bsdo = (int64_t)Lq * C; a.delta = delta; a.dq1 = dq; a.dk1 = dk; a.dv = dv; a.mode = 2;
    if (fod_attn_fwd(&a.fwd, nullptr)) { printf("fwd: %s\n", fod_last_error()); return 1; }
    for (int it = 0; it < 3; ++it)
        if (fod_attn_bwd(&a, nullptr)) { printf("bwd: %s\n", fod_last_error()); return 1; }
    std::vector<long long> tr(3 * 16 * 8);
    if (fod_debug_af_trace(tr.data())) { printf("trace read failed\n"); return 1; }
    const int nt = ((Lq + 127) / 128) * ((Lk + 127) / 128);
    long long t0 = tr[(0 * 16 + 0) * 8 + 0];
    printf("softmax thread 0 (clk since its first stamp): tile | wait s_full .. got | ld0 | exp0 done | g_done got | ld1(s_free) | stores done | p_full arrived\n");
    for (int t = 0; t < nt && t < 16; ++t) {
        printf("  t%-2d", t);
        for (int e = 0; e < 8; ++e) printf(" %7lld", tr[(0 * 16 + t) * 8 + e] - t0);
        printf("\n");
    }
    printf("MMA thread: tile | s_free(t-1) got | loads ready | S,dP issued | p_full got | grads issued\n");
    for (int t = 0; t < nt && t < 16; ++t) {
        printf("  t%-2d", t);
        for (int e = 0; e < 5; ++e) printf(" %7lld", tr[(1 * 16 + t) * 8 + e] ? tr[(1 * 16 + t) * 8 + e] - t0 : -1);
        printf("\n");
    }
    printf("CTA phases (clk since kernel entry): setup done | predecessor done | delta pre-pass done | main loop done | last MMAs retired | gradients stored | exit\n ");
    for (int e = 1; e < 8; ++e) printf(" %7lld", tr[(2 * 16 + 0) * 8 + e] - tr[(2 * 16 + 0) * 8 + 0]);
    printf("\n  (softmax first stamp at %lld)\n", t0 - tr[(2 * 16 + 0) * 8 + 0]);
    return 0;
}
