// AP bookkeeping of one forward pass: future_od/utils/od_map.py:214-287 prepare_od_map_stuffs as
// ONE launch — per (sample, class) CTA: top-50 detections by score, smoothed IoU against the annotation slots, and the
// 50-step greedy claim loop at 10 IoU thresholds (one warp per threshold, alive set as a per-lane bit mask), replacing
// ~330 small torch launches per step (sort, gather/scatter, max-reduce, where, compare in a 50-iteration host loop).
// Integer / boolean outputs are bit-exact; the float arithmetic follows torch's operation order (compiled -fmad=false).
// Latency-bound: ~60 KB of inputs per step.
#include "fod_internal.h"
#include "sm100_ptx.cuh"

namespace fod {

constexpr int AP_THREADS = 320;   // 10 warps = 10 IoU thresholds
constexpr int AP_MAX_M = 256;     // detections per sample
constexpr int AP_MAX_N = 256;     // annotation slots per sample
constexpr int AP_MAX_K = 64;      // top-k

struct ApArgs {
    const float* pred_boxes;   // (B, M, 4) xyxy pixels
    const float* scores;       // (B, M, C) sigmoid scores, generic class last
    const float* anno_boxes;   // (B, N, 4) xyxy pixels
    const long long* anno_classes;  // (B, N)
    const long long* anno_active;   // (B, N)
    const float* thresholds;   // (T)
    float* confs;              // (T, C, B*K)
    unsigned char* is_positive;     // (T, C, B*K) bool
    unsigned char* size_cat;        // (C, 4, B*K) bool
    long long* num_annos;      // (C, 4), zero-initialised by the caller
    int B, M, C, N, K, T;
    float size_lo, size_hi;    // area delimiters in pixels^2 (od_map.py:20-23 times H*W)
};

__device__ __forceinline__ float relu_f(float x) { return x > 0.f ? x : 0.f; }

__global__ void __launch_bounds__(AP_THREADS)
ap_match_kernel(const ApArgs a) {
    extern __shared__ float ap_smem[];
    float* iou_k = ap_smem;                                  // [K][N]
    __shared__ float s_score[AP_MAX_M];
    __shared__ int s_order[AP_MAX_K];
    __shared__ float s_conf[AP_MAX_K];
    __shared__ float s_pbox[AP_MAX_K][4];
    __shared__ unsigned char s_avail[AP_MAX_N];
    __shared__ int s_cnt[4];

    const int b = blockIdx.x, c = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int M = a.M, N = a.N, K = a.K, C = a.C;
    const long long bk = (long long)a.B * K;

    if (tid < 4) s_cnt[tid] = 0;
    for (int m = tid; m < M; m += AP_THREADS) s_score[m] = a.scores[((long long)b * M + m) * C + c];
    __syncthreads();

    // ---- top-K by descending score (rank by counting; equal scores keep ascending index order).  NaN sorts as the
    // largest value, as torch.sort does, so the ranks stay a permutation and every s_order slot is written.
    for (int m = tid; m < M; m += AP_THREADS) {
        const float s = s_score[m];
        const bool s_nan = s != s;
        int rank = 0;
        for (int j = 0; j < M; ++j) {
            const float t = s_score[j];
            const bool t_nan = t != t;
            const bool greater = (t_nan && !s_nan) || (!t_nan && !s_nan && t > s);
            const bool equal = (t_nan && s_nan) || (!t_nan && !s_nan && t == s);
            rank += greater || (equal && j < m);
        }
        if (rank < K) {
            s_order[rank] = m;
            s_conf[rank] = s;
        }
    }
    // ---- annotation availability for this class + annotation counts per size bucket (od_map.py:199-211)
    int cnt[4] = {0, 0, 0, 0};
    for (int n = tid; n < N; n += AP_THREADS) {
        const long long idx = (long long)b * N + n;
        const bool av = a.anno_active[idx] == 1 && (c == C - 1 || a.anno_classes[idx] == (long long)c);
        s_avail[n] = av ? 1 : 0;
        if (av) {
            const float* ab = a.anno_boxes + idx * 4;
            const float area = (ab[2] - ab[0]) * (ab[3] - ab[1]);
            cnt[0] += 1;
            cnt[1] += area <= a.size_lo;
            cnt[2] += (a.size_lo < area) && (area <= a.size_hi);
            cnt[3] += a.size_hi < area;
        }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        int v = cnt[s];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&s_cnt[s], v);
    }
    __syncthreads();
    if (tid < 4 && s_cnt[tid]) atomicAdd(reinterpret_cast<unsigned long long*>(a.num_annos + c * 4 + tid), (unsigned long long)s_cnt[tid]);

    // ---- per selected detection: box, confidence rows (same for every threshold), size buckets (od_map.py:163-181)
    for (int j = tid; j < K; j += AP_THREADS) {
        const float* pb = a.pred_boxes + ((long long)b * M + s_order[j]) * 4;
        const float x0 = pb[0], y0 = pb[1], x1 = pb[2], y1 = pb[3];
        s_pbox[j][0] = x0; s_pbox[j][1] = y0; s_pbox[j][2] = x1; s_pbox[j][3] = y1;
        const long long col = (long long)b * K + j;
        for (int t = 0; t < a.T; ++t) a.confs[((long long)t * C + c) * bk + col] = s_conf[j];
        const float area = (x1 - x0) * (y1 - y0);
        unsigned char* sc = a.size_cat + ((long long)c * 4) * bk + col;
        sc[0] = 1;
        sc[bk] = area <= a.size_lo;
        sc[2 * bk] = (a.size_lo < area) && (area <= a.size_hi);
        sc[3 * bk] = a.size_hi < area;
    }
    __syncthreads();

    // ---- smoothed IoU of the K detections against the N slots (od_map.py:48-71), zero where the slot is unavailable
    for (int n = tid; n < N; n += AP_THREADS) {
        const float* ab = a.anno_boxes + ((long long)b * N + n) * 4;
        const float ax0 = ab[0], ay0 = ab[1], ax1 = ab[2], ay1 = ab[3];
        const float area_a = relu_f(ax1 - ax0) * relu_f(ay1 - ay0);
        const bool av = s_avail[n] != 0;
        for (int j = 0; j < K; ++j) {
            const float px0 = s_pbox[j][0], py0 = s_pbox[j][1], px1 = s_pbox[j][2], py1 = s_pbox[j][3];
            const float area_p = relu_f(px1 - px0) * relu_f(py1 - py0);
            const float inter = relu_f(fminf(px1, ax1) - fmaxf(px0, ax0)) * relu_f(fminf(py1, ay1) - fmaxf(py0, ay0));
            const float iou = (inter + 1e-7f) / (((area_p + area_a) - inter) + 1e-7f);
            iou_k[j * N + n] = av ? iou : 0.f;
        }
    }
    __syncthreads();

    // ---- greedy claim loop (od_map.py:267-277): warp t handles threshold t; lane l owns slots l, l+32, ...
    if (warp < a.T) {
        const float thr = a.thresholds[warp];
        unsigned alive = 0xffffffffu;  // bit i: slot lane + 32*i still unclaimed (N <= 32*32)
        unsigned char* pos = a.is_positive + ((long long)warp * C + c) * bk + (long long)b * K;
        for (int j = 0; j < K; ++j) {
            float best = -1.f;
            int arg = 0x7fffffff;
            const float* row = iou_k + j * N;
            for (int i = 0, n = lane; n < N; ++i, n += 32) {
                const float v = ((alive >> i) & 1u) ? row[n] : 0.f;  // iou * alive: claimed slots still compete with 0
                if (v > best) {                                        // ascending n: first maximum wins
                    best = v;
                    arg = n;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
                if (ob > best || (ob == best && oa < arg)) {
                    best = ob;
                    arg = oa;
                }
            }
            const bool hit = best >= thr;
            if (hit && (arg & 31) == lane) alive &= ~(1u << (arg >> 5));
            if (lane == 0) pos[j] = hit ? 1 : 0;
        }
    }
}

}  // namespace fod

extern "C" int fod_ap_match(const fod_ap_args* a, void* stream) {
    using namespace fod;
    if (!a || !a->pred_boxes || !a->scores || !a->anno_boxes || !a->anno_classes || !a->anno_active || !a->thresholds ||
        !a->confs || !a->is_positive || !a->size_cat || !a->num_annos)
        return fod_fail(FOD_E_ARG, "ap_match: null argument");
    if (a->B <= 0 || a->M <= 0 || a->C <= 0 || a->N <= 0 || a->K <= 0 || a->T <= 0)
        return fod_fail(FOD_E_ARG, "ap_match: empty");
    if (a->M > AP_MAX_M || a->N > AP_MAX_N || a->K > AP_MAX_K || a->K > a->M || a->T > AP_THREADS / 32)
        return fod_fail(FOD_E_ARG, "ap_match: M,N <= 256, K <= min(64, M), T <= 10");
    ApArgs k;
    k.pred_boxes = a->pred_boxes; k.scores = a->scores; k.anno_boxes = a->anno_boxes;
    k.anno_classes = reinterpret_cast<const long long*>(a->anno_classes);
    k.anno_active = reinterpret_cast<const long long*>(a->anno_active);
    k.thresholds = a->thresholds; k.confs = a->confs;
    k.is_positive = reinterpret_cast<unsigned char*>(a->is_positive);
    k.size_cat = reinterpret_cast<unsigned char*>(a->size_cat);
    k.num_annos = reinterpret_cast<long long*>(a->num_annos);
    k.B = a->B; k.M = a->M; k.C = a->C; k.N = a->N; k.K = a->K; k.T = a->T;
    k.size_lo = a->size_lo; k.size_hi = a->size_hi;
    const int smem = a->K * a->N * (int)sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(ap_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_MAX_K * AP_MAX_N * 4);
        attr_set = true;
    }
    dim3 grid(a->B, a->C);
    ap_match_kernel<<<grid, AP_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(k);
    return fod_check_launch("ap_match_kernel");
}
