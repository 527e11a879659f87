// Standalone device self-test for the tcgen05 kernels (no Python): compares fod_linear_fwd / fod_attn_fwd
// against naive SIMT kernels on the same bf16 inputs and prints max errors + CUDA-event timings.
// Build: make selftest ; run on a B200: ./selftest   (exit code 0 = all within tolerance)
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../include/fod_b200.h"

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e_ = (x);                                                              \
        if (e_ != cudaSuccess) {                                                           \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                       \
        }                                                                                  \
    } while (0)

static uint32_t g_seed = 12345;
static float frand() {
    g_seed = g_seed * 1664525u + 1013904223u;
    return ((g_seed >> 8) & 0xFFFF) / 32768.f - 1.f;
}
static __nv_bfloat16* dev_bf16(size_t n, float scale, std::vector<float>* keep = nullptr) {
    std::vector<__nv_bfloat16> h(n);
    if (keep) keep->resize(n);
    for (size_t i = 0; i < n; ++i) {
        h[i] = __float2bfloat16(frand() * scale);
        if (keep) (*keep)[i] = __bfloat162float(h[i]);
    }
    __nv_bfloat16* d;
    CK(cudaMalloc(&d, n * 2));
    CK(cudaMemcpy(d, h.data(), n * 2, cudaMemcpyHostToDevice));
    return d;
}
static float* dev_f32(size_t n, float scale, float offset = 0.f) {
    std::vector<float> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = frand() * scale + offset;
    float* d;
    CK(cudaMalloc(&d, n * 4));
    CK(cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice));
    return d;
}

__global__ void ref_linear(const __nv_bfloat16* x, const __nv_bfloat16* w, const float* bias,
                           const __nv_bfloat16* res, int res_mod, int relu, float* y, int M, int N, int K) {
    int n = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y;
    if (n >= N) return;
    float acc = 0.f;
    for (int k = 0; k < K; ++k) acc += __bfloat162float(x[(size_t)m * K + k]) * __bfloat162float(w[(size_t)n * K + k]);
    if (bias) acc += bias[n];
    if (res) acc += __bfloat162float(res[(size_t)(res_mod ? m % res_mod : m) * N + n]);
    if (relu) acc = fmaxf(acc, 0.f);
    y[(size_t)m * N + n] = acc;
}
__global__ void ref_ln(float* y, const float* g, const float* b, int M, int N, float eps) {
    int m = blockIdx.x;
    if (threadIdx.x) return;
    float s = 0;
    for (int i = 0; i < N; ++i) s += y[(size_t)m * N + i];
    float mean = s / N, v = 0;
    for (int i = 0; i < N; ++i) {
        float d = y[(size_t)m * N + i] - mean;
        v += d * d;
    }
    float r = rsqrtf(v / N + eps);
    for (int i = 0; i < N; ++i) y[(size_t)m * N + i] = (y[(size_t)m * N + i] - mean) * r * g[i] + b[i];
}
// one thread per (b,h,q): full softmax row
__global__ void ref_attn(const __nv_bfloat16* q1, const __nv_bfloat16* k1, const __nv_bfloat16* q2,
                         const __nv_bfloat16* k2, const __nv_bfloat16* v, float* out, float* lse, int B, int H,
                         int Lq, int Lk, float scale, int k2_bcast) {
    int q = blockIdx.x * blockDim.x + threadIdx.x, h = blockIdx.y, b = blockIdx.z;
    if (q >= Lq) return;
    const int C = H * 32;
    float mx = -INFINITY;
    for (int pass = 0; pass < 2; ++pass) {
        float l = 0.f, o[32];
        for (int i = 0; i < 32; ++i) o[i] = 0.f;
        for (int j = 0; j < Lk; ++j) {
            float s = 0.f;
            for (int d = 0; d < 32; ++d) {
                s += __bfloat162float(q1[((size_t)b * Lq + q) * C + h * 32 + d]) *
                     __bfloat162float(k1[((size_t)b * Lk + j) * C + h * 32 + d]);
                if (q2)
                    s += __bfloat162float(q2[((size_t)b * Lq + q) * C + h * 32 + d]) *
                         __bfloat162float(k2[((size_t)(k2_bcast ? 0 : b) * Lk + j) * C + h * 32 + d]);
            }
            s *= scale;
            if (pass == 0) {
                mx = fmaxf(mx, s);
            } else {
                float pj = expf(s - mx);
                l += pj;
                for (int d = 0; d < 32; ++d) o[d] += pj * __bfloat162float(v[((size_t)b * Lk + j) * C + h * 32 + d]);
            }
        }
        if (pass == 1) {
            for (int d = 0; d < 32; ++d) out[((size_t)b * Lq + q) * C + h * 32 + d] = o[d] / l;
            lse[((size_t)b * H + h) * Lq + q] = mx + logf(l);
        }
    }
}

static int g_fail = 0;
static void report(const char* name, double err, double tol, float ms, double flops) {
    printf("%-44s max_err %.4g (tol %.3g) %s  %.3f ms  %.1f TFLOP/s\n", name, err, tol, err <= tol ? "PASS" : "FAIL", ms,
           flops / (ms * 1e9));
    if (!(err <= tol)) g_fail = 1;
}

static void test_linear(int M, int N, int K, int bias, int res, int res_mod, int relu, int ln, int f32out, int save_pre = 0) {
    __nv_bfloat16* x = dev_bf16((size_t)M * K, 1.f);
    __nv_bfloat16* w = dev_bf16((size_t)N * K, 0.08f);
    float* b = bias ? dev_f32(N, 0.5f) : nullptr;
    int res_rows = res_mod ? res_mod : M;
    __nv_bfloat16* r = res ? dev_bf16((size_t)res_rows * N, 1.f) : nullptr;
    float* g = ln ? dev_f32(N, 0.3f, 1.f) : nullptr;
    float* be = ln ? dev_f32(N, 0.3f) : nullptr;
    float *yref, *yf32 = nullptr, *mean = nullptr, *rstd = nullptr;
    __nv_bfloat16* y;
    CK(cudaMalloc(&yref, (size_t)M * N * 4));
    CK(cudaMalloc(&y, (size_t)M * N * 2));
    CK(cudaMemset(y, 0xff, (size_t)M * N * 2));
    if (f32out) CK(cudaMalloc(&yf32, (size_t)M * N * 4));
    __nv_bfloat16* pre = nullptr;
    if (ln) {
        CK(cudaMalloc(&mean, M * 4));
        CK(cudaMalloc(&rstd, M * 4));
        if (save_pre) CK(cudaMalloc(&pre, (size_t)M * N * 2));
    }
    ref_linear<<<dim3((N + 127) / 128, M), 128>>>(x, w, b, r, res_mod, relu, yref, M, N, K);
    if (ln) ref_ln<<<M, 32>>>(yref, g, be, M, N, 1e-5f);
    CK(cudaDeviceSynchronize());
    fod_linear_args a = {};
    a.x = x; a.w = w; a.bias = b; a.residual = r; a.ln_gamma = g; a.ln_beta = be;
    a.out_bf16 = f32out ? nullptr : y; a.out_f32 = yf32; a.ln_mean = mean; a.ln_rstd = rstd;
    a.ldx = K; a.ld_res = N; a.ld_out = N; a.M = M; a.N = N; a.K = K; a.res_row_mod = res_mod; a.relu = relu;
    a.ln_eps = 1e-5f;
    a.pre_ln_bf16 = pre;
    int rc = fod_linear_fwd(&a, nullptr);
    if (rc) { printf("fod_linear_fwd rc=%d (%s)\n", rc, fod_last_error()); g_fail = 1; return; }
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) fod_linear_fwd(&a, nullptr);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
    std::vector<float> hr((size_t)M * N);
    CK(cudaMemcpy(hr.data(), yref, hr.size() * 4, cudaMemcpyDeviceToHost));
    double err = 0;
    if (f32out) {
        std::vector<float> hy(hr.size());
        CK(cudaMemcpy(hy.data(), yf32, hy.size() * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < hr.size(); ++i) err = fmax(err, fabs(hy[i] - hr[i]) / (1.0 + fabs(hr[i])));
    } else {
        std::vector<__nv_bfloat16> hy(hr.size());
        CK(cudaMemcpy(hy.data(), y, hy.size() * 2, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < hr.size(); ++i) {
            double d = fabs(__bfloat162float(hy[i]) - hr[i]) / (1.0 + fabs(hr[i]));
            if (!(d == d)) d = 1e9;
            err = fmax(err, d);
        }
    }
    char name[128];
    snprintf(name, sizeof name, "linear M%d N%d K%d b%d r%d/%d relu%d ln%d%s f32%d", M, N, K, bias, res, res_mod, relu, ln, save_pre ? "+pre" : "", f32out);
    report(name, err, f32out ? 2e-3 : 1.2e-2, ms, 2.0 * M * N * K);
    cudaFree(x); cudaFree(w); cudaFree(b); cudaFree(r); cudaFree(g); cudaFree(be); cudaFree(yref); cudaFree(y);
    cudaFree(yf32); cudaFree(mean); cudaFree(rstd); cudaFree(pre);
}

// keys scaled by a ramp over the key index: row maxima keep growing from one key tile to the next, which drives the
// forward kernel through its accumulator-rescaling path on (almost) every tile
__global__ void ramp_keys(__nv_bfloat16* k, int Lk, int C, float ramp, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int key = int((i / C) % Lk);
    k[i] = __float2bfloat16(__bfloat162float(k[i]) * (1.f + ramp * key / Lk));
}

static void test_attn(int B, int H, int Lq, int Lk, int two, int k2_bcast, float ramp = 0.f) {
    const int C = H * 32;
    __nv_bfloat16* q1 = dev_bf16((size_t)B * Lq * C, 1.5f);
    __nv_bfloat16* k1 = dev_bf16((size_t)B * Lk * C, 1.5f);
    if (ramp != 0.f) {
        const size_t n = (size_t)B * Lk * C;
        ramp_keys<<<(unsigned)((n + 255) / 256), 256>>>(k1, Lk, C, ramp, n);
        CK(cudaDeviceSynchronize());
    }
    __nv_bfloat16* q2 = two ? dev_bf16((size_t)B * Lq * C, 1.5f) : nullptr;
    __nv_bfloat16* k2 = two ? dev_bf16((size_t)(k2_bcast ? 1 : B) * Lk * C, 1.5f) : nullptr;
    __nv_bfloat16* v = dev_bf16((size_t)B * Lk * C, 1.f);
    float *oref, *lref, *lse;
    __nv_bfloat16* o;
    CK(cudaMalloc(&oref, (size_t)B * Lq * C * 4));
    CK(cudaMalloc(&lref, (size_t)B * H * Lq * 4));
    CK(cudaMalloc(&lse, (size_t)B * H * Lq * 4));
    CK(cudaMalloc(&o, (size_t)B * Lq * C * 2));
    CK(cudaMemset(o, 0xff, (size_t)B * Lq * C * 2));
    const float scale = two ? 0.125f : 0.17677669f;
    ref_attn<<<dim3((Lq + 63) / 64, H, B), 64>>>(q1, k1, q2, k2, v, oref, lref, B, H, Lq, Lk, scale, k2_bcast);
    CK(cudaDeviceSynchronize());
    fod_attn_args a = {};
    a.q1 = q1; a.k1 = k1; a.q2 = q2; a.k2 = k2; a.v = v; a.out = o; a.lse = lse;
    a.ldq1 = a.ldk1 = a.ldq2 = a.ldk2 = a.ldv = a.ldo = C;
    a.bsq1 = a.bsq2 = a.bso = (int64_t)Lq * C;
    a.bsk1 = a.bsv = (int64_t)Lk * C;
    a.bsk2 = k2_bcast ? 0 : (int64_t)Lk * C;
    a.B = B; a.H = H; a.Lq = Lq; a.Lk = Lk; a.scale = scale;
    int rc = fod_attn_fwd(&a, nullptr);
    if (rc) { printf("fod_attn_fwd rc=%d (%s)\n", rc, fod_last_error()); g_fail = 1; return; }
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) fod_attn_fwd(&a, nullptr);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
    std::vector<float> hr((size_t)B * Lq * C), hl((size_t)B * H * Lq), hl2(hl.size());
    std::vector<__nv_bfloat16> hy(hr.size());
    CK(cudaMemcpy(hr.data(), oref, hr.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hy.data(), o, hy.size() * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hl.data(), lref, hl.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hl2.data(), lse, hl.size() * 4, cudaMemcpyDeviceToHost));
    double err = 0, lerr = 0;
    for (size_t i = 0; i < hr.size(); ++i) {
        double d = fabs(__bfloat162float(hy[i]) - hr[i]);
        if (!(d == d)) d = 1e9;
        err = fmax(err, d);
    }
    for (size_t i = 0; i < hl.size(); ++i) lerr = fmax(lerr, fabs(hl[i] - hl2[i]));
    char name[128];
    snprintf(name, sizeof name, "attn B%d H%d Lq%d Lk%d parts%d bc%d ramp%g (lse err %.3g)", B, H, Lq, Lk, two + 1, k2_bcast, ramp, lerr);
    report(name, fmax(err, lerr), 2e-2, ms, (two ? 6.0 : 4.0) * B * H * (double)Lq * Lk * 32);
    cudaFree(q1); cudaFree(k1); cudaFree(q2); cudaFree(k2); cudaFree(v); cudaFree(oref); cudaFree(lref);
    cudaFree(lse); cudaFree(o);
}

// Attention backward: the fused single kernel (mode 2) against the two-kernel path (mode 1) on the same inputs, plus
// timings of both.  The two paths share no code below the argument checks, so agreement to bf16 rounding pins both.
static void test_attn_bwd(int B, int H, int Lq, int Lk, int two, int k2_bcast) {
    const int C = H * 32;
    __nv_bfloat16* q1 = dev_bf16((size_t)B * Lq * C, 1.5f);
    __nv_bfloat16* k1 = dev_bf16((size_t)B * Lk * C, 1.5f);
    __nv_bfloat16* q2 = two ? dev_bf16((size_t)B * Lq * C, 1.5f) : nullptr;
    __nv_bfloat16* k2 = two ? dev_bf16((size_t)(k2_bcast ? 1 : B) * Lk * C, 1.5f) : nullptr;
    __nv_bfloat16* v = dev_bf16((size_t)B * Lk * C, 1.f);
    __nv_bfloat16* dout = dev_bf16((size_t)B * Lq * C, 1.f);
    __nv_bfloat16* o;
    float *lse, *delta[2];
    CK(cudaMalloc(&o, (size_t)B * Lq * C * 2));
    CK(cudaMalloc(&lse, (size_t)B * H * Lq * 4));
    const size_t nq = (size_t)B * Lq * C, nk = (size_t)B * Lk * C;
    __nv_bfloat16* g[2][5];  // dq1, dq2, dk1, dk2, dv per mode
    for (int m = 0; m < 2; ++m) {
        CK(cudaMalloc(&delta[m], (size_t)B * H * Lq * 4));
        for (int t = 0; t < 5; ++t) {
            const size_t n = (t < 2 ? nq : nk);
            CK(cudaMalloc(&g[m][t], n * 2));
            CK(cudaMemset(g[m][t], 0xff, n * 2));
        }
    }
    const float scale = two ? 0.125f : 0.17677669f;
    fod_attn_bwd_args a = {};
    a.fwd.q1 = q1; a.fwd.k1 = k1; a.fwd.q2 = q2; a.fwd.k2 = k2; a.fwd.v = v; a.fwd.out = o; a.fwd.lse = lse;
    a.fwd.ldq1 = a.fwd.ldk1 = a.fwd.ldq2 = a.fwd.ldk2 = a.fwd.ldv = a.fwd.ldo = C;
    a.fwd.bsq1 = a.fwd.bsq2 = a.fwd.bso = (int64_t)Lq * C;
    a.fwd.bsk1 = a.fwd.bsv = (int64_t)Lk * C;
    a.fwd.bsk2 = k2_bcast ? 0 : (int64_t)Lk * C;
    a.fwd.B = B; a.fwd.H = H; a.fwd.Lq = Lq; a.fwd.Lk = Lk; a.fwd.scale = scale;
    a.dout = dout; a.lddo = C; a.bsdo = (int64_t)Lq * C;
    int rc = fod_attn_fwd(&a.fwd, nullptr);
    if (rc) { printf("fod_attn_fwd rc=%d (%s)\n", rc, fod_last_error()); g_fail = 1; return; }
    float ms[2] = {0, 0};
    for (int m = 0; m < 2; ++m) {
        a.mode = m + 1;
        a.delta = delta[m];
        a.dq1 = g[m][0]; a.dq2 = two ? g[m][1] : nullptr; a.dk1 = g[m][2]; a.dk2 = two ? g[m][3] : nullptr; a.dv = g[m][4];
        rc = fod_attn_bwd(&a, nullptr);
        if (rc) { printf("fod_attn_bwd mode %d rc=%d (%s)\n", m + 1, rc, fod_last_error()); g_fail = 1; return; }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("attn_bwd mode %d: %s\n", m + 1, cudaGetErrorString(e)); exit(2); }
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        for (int i = 0; i < 10; ++i) fod_attn_bwd(&a, nullptr);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&ms[m], e0, e1); ms[m] /= 10;
    }
    double worst = 0;
    const char* names[5] = {"dq1", "dq2", "dk1", "dk2", "dv"};
    char detail[256] = "";
    for (int t = 0; t < 5; ++t) {
        if (!two && (t == 1 || t == 3)) continue;
        const size_t n = (t < 2 ? nq : nk);
        std::vector<__nv_bfloat16> h0(n), h1(n);
        CK(cudaMemcpy(h0.data(), g[0][t], n * 2, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(h1.data(), g[1][t], n * 2, cudaMemcpyDeviceToHost));
        double num = 0, den = 0;
        for (size_t i = 0; i < n; ++i) {
            const double x = __bfloat162float(h0[i]), y = __bfloat162float(h1[i]);
            double d = y - x;
            if (!(d == d)) d = 1e6;
            num += d * d; den += x * x;
        }
        const double rel = sqrt(num / (den > 0 ? den : 1));
        if (rel > 1e-2) {  // where do the two paths part?  first rows of both, and the first differing element
            printf("  %s two-kernel:", names[t]);
            for (int i = 0; i < 8; ++i) printf(" %.4f", __bfloat162float(h0[i]));
            printf("\n  %s fused     :", names[t]);
            for (int i = 0; i < 8; ++i) printf(" %.4f", __bfloat162float(h1[i]));
            size_t first = n, cnt = 0;
            for (size_t i = 0; i < n; ++i) {
                const float x = __bfloat162float(h0[i]), y = __bfloat162float(h1[i]);
                if (!(fabsf(x - y) <= 0.02f * (fabsf(x) + 0.05f))) { if (first == n) first = i; ++cnt; }
            }
            const size_t Cc = (size_t)H * 32;
            printf("\n  %zu of %zu differ; first at row %zu col %zu\n", cnt, n, first / Cc, first % Cc);
        }
        worst = fmax(worst, rel);
        snprintf(detail + strlen(detail), sizeof(detail) - strlen(detail), " %s %.2g", names[t], rel);
    }
    {
        std::vector<float> d0((size_t)B * H * Lq), d1(d0.size());
        CK(cudaMemcpy(d0.data(), delta[0], d0.size() * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(d1.data(), delta[1], d1.size() * 4, cudaMemcpyDeviceToHost));
        double dm = 0;
        for (size_t i = 0; i < d0.size(); ++i) dm = fmax(dm, fabs(d0[i] - d1[i]));
        snprintf(detail + strlen(detail), sizeof(detail) - strlen(detail), " delta %.2g", dm);
        if (dm > 1e-3) worst = fmax(worst, 1.0);
    }
    char name[320];
    snprintf(name, sizeof name, "attn_bwd B%d H%d Lq%d Lk%d parts%d bc%d two-kernel %.3f ms | fused vs two-kernel rel-L2:%s", B, H, Lq, Lk,
             two + 1, k2_bcast, ms[0], detail);
    report(name, worst, 1e-2, ms[1], (two ? 16.0 : 10.0) * B * H * (double)Lq * Lk * 32);
    cudaFree(q1); cudaFree(k1); cudaFree(q2); cudaFree(k2); cudaFree(v); cudaFree(dout); cudaFree(o); cudaFree(lse);
    for (int m = 0; m < 2; ++m) { cudaFree(delta[m]); for (int t = 0; t < 5; ++t) cudaFree(g[m][t]); }
}

int main(int argc, char** argv) {
    int which = argc > 1 ? atoi(argv[1]) : 3;
    printf("fod_b200 selftest, abi %d\n", fod_version());
    if (which & 1) {
        test_linear(256, 256, 64, 0, 0, 0, 0, 0, 0);
        test_linear(256, 256, 256, 1, 0, 0, 0, 0, 0);
        test_linear(5600, 2048, 256, 1, 0, 0, 1, 0, 0);
        test_linear(5600, 256, 2048, 1, 1, 0, 0, 1, 0);
        test_linear(5600, 768, 256, 1, 1, 350, 0, 0, 0);
        test_linear(1000, 8, 256, 1, 0, 0, 0, 0, 1);
        test_linear(1024, 128, 256, 1, 0, 0, 1, 0, 0);
        test_linear(22400, 2048, 256, 1, 0, 0, 1, 0, 0);
        test_linear(22400, 256, 2048, 1, 1, 0, 0, 1, 0);
    }
    if (which & 4) {  // the LayerNorm-epilogue GEMMs of the training step (pre-LN sum, mean, rstd saved for backward)
        test_linear(1024, 256, 256, 1, 1, 0, 0, 1, 0, 1);
        test_linear(1024, 256, 2048, 1, 1, 0, 0, 1, 0, 1);
        test_linear(4096, 256, 256, 1, 1, 0, 0, 1, 0, 1);
        test_linear(4096, 256, 2048, 1, 1, 0, 0, 1, 0, 1);
        test_linear(22400, 256, 256, 1, 1, 0, 0, 1, 0, 1);
        test_linear(22400, 256, 2048, 1, 1, 0, 0, 1, 0, 1);
        test_linear(22400, 768, 256, 1, 1, 350, 0, 0, 0);
        test_linear(22400, 2048, 256, 1, 0, 0, 1, 0, 0);
        test_linear(4096, 768, 256, 1, 1, 128, 0, 0, 0);
    }
    if (which & 16) test_attn_bwd(2, 8, 128, 350, 1, 1);  // single cases (a device fault ends the process)
    if (which & 32) test_attn_bwd(2, 8, 128, 350, 1, 0);
    if (which & 64) test_attn_bwd(2, 8, 128, 350, 0, 0);
    if (which & 128) test_attn_bwd(2, 8, 128, 128, 1, 1);
    if (which & 256) test_attn_bwd(1, 1, 128, 128, 1, 0);
    if (which & 8) {
        test_attn_bwd(1, 1, 128, 128, 0, 0);
        test_attn_bwd(2, 8, 350, 350, 0, 0);
        test_attn_bwd(2, 8, 128, 350, 1, 1);
        test_attn_bwd(2, 8, 200, 300, 1, 0);
        test_attn_bwd(3, 8, 768, 129, 0, 0);
        test_attn_bwd(16, 8, 350, 350, 0, 0);
        test_attn_bwd(64, 8, 350, 350, 0, 0);
        test_attn_bwd(32, 8, 128, 350, 1, 1);
        test_attn_bwd(32, 8, 128, 128, 0, 0);
        test_attn_bwd(16, 8, 700, 700, 0, 0);
    }
    if (which & 2) {
        test_attn(1, 1, 128, 128, 0, 0);
        test_attn(2, 8, 350, 350, 0, 0);
        test_attn(2, 8, 128, 350, 1, 1);
        test_attn(2, 8, 128, 128, 0, 0);
        test_attn(16, 8, 350, 350, 0, 0);
        test_attn(16, 8, 1400, 1400, 0, 0);
        test_attn(2, 8, 2800, 2800, 0, 0);
        test_attn(2, 8, 350, 350, 0, 0, 40.f);
        test_attn(2, 8, 128, 350, 1, 1, 40.f);
        test_attn(4, 8, 1400, 1400, 0, 0, 60.f);
        test_attn(4, 8, 1400, 1400, 0, 0, -0.9f);
    }
    printf(g_fail ? "SELFTEST FAILED\n" : "SELFTEST OK\n");
    return g_fail;
}
