// Set-prediction loss path on the device (north-star piece 4):
//   fod_targets_compact : st_detr.py:237-263 (to_detr_targets) without host syncs
//   fod_match           : ConditionalDETR/models/matcher.py:58-86 — per-sample cost block (focal class cost,
//                         L1 cdist, -GIoU) in fp32 with torch's operation order (no FMA contraction: this file is
//                         compiled with -fmad=false), then the rectangular shortest-augmenting-path solver of
//                         scipy.optimize.linear_sum_assignment (Crouse 2016) in fp64, one warp per (level, sample),
//                         warp-shuffle arg-min with scipy's scan-order tie rule.
//   fod_set_loss        : set_criterion.py:36-115 + segmentation.py:198-223 — focal / L1 / GIoU losses, their
//                         gradients w.r.t. logits and boxes, cardinality and class error, from the dense assignment.
// All CUDA-core work: latency-bound (a few hundred KB per step); the point is removing 6 D2H syncs + host scipy.
#include <math.h>
#include "fod_internal.h"

namespace fod {

constexpr int MT_THREADS = 128;

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// ------------------------------------------------------------------------------------------------ targets
__global__ void targets_compact_kernel(const float* __restrict__ boxes_xyxy, const int64_t* __restrict__ classes,
                                       const int64_t* __restrict__ active, float* __restrict__ out_boxes,
                                       int32_t* __restrict__ out_labels, int32_t* __restrict__ out_count,
                                       float* __restrict__ total_count, int slots, float inv_w, float inv_h) {
    const int b = blockIdx.x;
    const int lane = threadIdx.x;  // one warp
    int base = 0;
    for (int s0 = 0; s0 < slots; s0 += 32) {
        const int s = s0 + lane;
        const bool on = s < slots && active[(long long)b * slots + s] == 1;
        const unsigned m = __ballot_sync(0xffffffffu, on);
        if (on) {
            const int dst = base + __popc(m & ((1u << lane) - 1u));
            const float* bx = boxes_xyxy + ((long long)b * slots + s) * 4;
            float* o = out_boxes + ((long long)b * slots + dst) * 4;
            o[0] = (0.5f * (bx[0] + bx[2])) * inv_w;
            o[1] = (0.5f * (bx[1] + bx[3])) * inv_h;
            o[2] = (bx[2] - bx[0]) * inv_w;
            o[3] = (bx[3] - bx[1]) * inv_h;
            out_labels[(long long)b * slots + dst] = (int32_t)classes[(long long)b * slots + s];
        }
        base += __popc(m);
    }
    if (lane == 0) {
        out_count[b] = base;
        if (total_count) atomicAdd(total_count, (float)base);  // integer-valued floats: order-independent
    }
}

// ------------------------------------------------------------------------------------------------ matcher
struct MatchParams {
    const float* logits;
    const float* boxes;
    const float* tgt_boxes;
    const int32_t* tgt_labels;
    const int32_t* tgt_count;
    int32_t* match_tgt;
    int64_t* pred_idx;
    int64_t* tgt_idx;
    int32_t* status;
    float* cost_out;
    float* cost_ws;  // used when the cost block does not fit shared memory
    int L, B, M, C, NT, KMAX;
    int smem_cost_floats;
    float w_class, w_bbox, w_giou, alpha;
};

struct Cand {
    double d;
    int t;
    int a;
};
// scipy's scan rule as an order-independent combine: smaller distance wins; among equal distances the LAST
// unassigned column in scan order wins, else the FIRST column.
__device__ __forceinline__ Cand cand_combine(const Cand& x, const Cand& y) {
    if (x.t < 0) return y;
    if (y.t < 0) return x;
    if (x.d < y.d) return x;
    if (y.d < x.d) return y;
    if (x.a && y.a) return x.t > y.t ? x : y;
    if (x.a) return x;
    if (y.a) return y;
    return x.t < y.t ? x : y;
}

__global__ void __launch_bounds__(MT_THREADS) match_kernel(const MatchParams p) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int lb = blockIdx.x;  // level * B + b
    const int b = lb % p.B;
    const int M = p.M, C = p.C;
    int nt = p.tgt_count[b];
    if (nt > p.NT) nt = p.NT;
    const int tid = threadIdx.x;
    const bool transposed = nt < M;  // scipy solves the transpose when rows > cols
    const int nr = transposed ? nt : M;
    const int nc = transposed ? M : nt;
    const int ncmax = max(M, p.NT), nrmax = min(M, p.NT);

    // shared carve-up
    double* u = reinterpret_cast<double*>(sm_raw);
    double* v = u + nrmax;
    double* dist = v + ncmax;
    int* pred = reinterpret_cast<int*>(dist + ncmax);
    int* row4col = pred + ncmax;
    int* col4row = row4col + ncmax;
    int* todo = col4row + nrmax;
    unsigned char* in_sr = reinterpret_cast<unsigned char*>(todo + ncmax);
    unsigned char* in_sc = in_sr + nrmax;
    uintptr_t off = reinterpret_cast<uintptr_t>(in_sc + ncmax);
    off = (off + 15) & ~uintptr_t(15);
    float* cc = reinterpret_cast<float*>(off);  // (M, C) class cost pos - neg
    float* cost_sm = cc + M * C;
    float* cost = ((long long)M * nt <= p.smem_cost_floats) ? cost_sm : p.cost_ws + (long long)lb * M * p.NT;
    __shared__ int s_bad;
    if (tid == 0) s_bad = 0;

    int32_t* mt = p.match_tgt + (long long)lb * M;
    for (int q = tid; q < M; q += MT_THREADS) mt[q] = -1;
    if (p.pred_idx)
        for (int k = tid; k < p.KMAX; k += MT_THREADS) {
            p.pred_idx[(long long)lb * p.KMAX + k] = -1;
            p.tgt_idx[(long long)lb * p.KMAX + k] = -1;
        }
    if (nt <= 0) return;

    const float* lg = p.logits + (long long)lb * M * C;
    const float* bx = p.boxes + (long long)lb * M * 4;
    const float* tb = p.tgt_boxes + (long long)b * p.NT * 4;
    const int32_t* tl = p.tgt_labels + (long long)b * p.NT;

    // class cost per (query, class): matcher.py:69-73
    for (int e = tid; e < M * C; e += MT_THREADS) {
        const float pr = sigmoidf_(lg[e]);
        const float neg = ((1.f - p.alpha) * (pr * pr)) * (-logf((1.f - pr) + 1e-8f));
        const float om = 1.f - pr;
        const float pos = (p.alpha * (om * om)) * (-logf(pr + 1e-8f));
        cc[e] = pos - neg;
    }
    __syncthreads();
    int bad = 0;
    for (int e = tid; e < M * nt; e += MT_THREADS) {
        const int q = e / nt, t = e - q * nt;
        const float qcx = bx[q * 4], qcy = bx[q * 4 + 1], qw = bx[q * 4 + 2], qh = bx[q * 4 + 3];
        const float tcx = tb[t * 4], tcy = tb[t * 4 + 1], tw = tb[t * 4 + 2], th = tb[t * 4 + 3];
        const float l1 = ((fabsf(qcx - tcx) + fabsf(qcy - tcy)) + fabsf(qw - tw)) + fabsf(qh - th);
        const float qx0 = qcx - 0.5f * qw, qy0 = qcy - 0.5f * qh, qx1 = qcx + 0.5f * qw, qy1 = qcy + 0.5f * qh;
        const float tx0 = tcx - 0.5f * tw, ty0 = tcy - 0.5f * th, tx1 = tcx + 0.5f * tw, ty1 = tcy + 0.5f * th;
        if (!(qx1 >= qx0) || !(qy1 >= qy0) || !(tx1 >= tx0) || !(ty1 >= ty0)) bad |= 2;  // box_ops.py:59-60
        const float aq = (qx1 - qx0) * (qy1 - qy0), at = (tx1 - tx0) * (ty1 - ty0);
        const float iw = fmaxf(fminf(qx1, tx1) - fmaxf(qx0, tx0), 0.f);
        const float ih = fmaxf(fminf(qy1, ty1) - fmaxf(qy0, ty0), 0.f);
        const float inter = iw * ih;
        const float uni = (aq + at) - inter;
        const float iou = inter / uni;
        const float hw = fmaxf(fmaxf(qx1, tx1) - fminf(qx0, tx0), 0.f);
        const float hh = fmaxf(fmaxf(qy1, ty1) - fminf(qy0, ty0), 0.f);
        const float hull = hw * hh;
        const float giou = iou - (hull - uni) / hull;
        int lab = tl[t];
        lab = lab < 0 ? 0 : (lab >= C ? C - 1 : lab);
        const float cst = (p.w_bbox * l1 + p.w_class * cc[q * C + lab]) + p.w_giou * (-giou);
        if (!(cst == cst) || cst == -INFINITY) bad |= 1;  // scipy: "matrix contains invalid numeric entries"
        if (p.cost_out) p.cost_out[((long long)lb * M + q) * p.NT + t] = cst;
        cost[transposed ? t * M + q : q * nt + t] = cst;
    }
    if (bad) atomicOr(&s_bad, bad);
    __syncthreads();
    if (s_bad) {
        if (tid == 0) atomicOr(p.status, s_bad);
        if (s_bad & 1) return;
    }
    if (tid >= 32) return;

    // ---------------- LSAP: warp 0, scipy rectangular_lsap.cpp semantics (see oracle/lsap.c)
    const int lane = tid;
    const unsigned FULL = 0xffffffffu;
    for (int i = lane; i < nr; i += 32) { u[i] = 0.0; col4row[i] = -1; }
    for (int j = lane; j < nc; j += 32) { v[j] = 0.0; row4col[j] = -1; pred[j] = -1; }
    __syncwarp();
    for (int cur = 0; cur < nr; ++cur) {
        for (int i = lane; i < nr; i += 32) in_sr[i] = 0;
        for (int j = lane; j < nc; j += 32) { in_sc[j] = 0; dist[j] = INFINITY; todo[j] = nc - 1 - j; }
        __syncwarp();
        double lo = 0.0;
        int n_todo = nc, i = cur, sink = -1;
        while (sink < 0) {
            if (lane == 0) in_sr[i] = 1;
            const double ui = u[i];
            const float* crow = cost + (long long)i * nc;
            Cand best;
            best.d = INFINITY; best.t = -1; best.a = 0;
            for (int t = lane; t < n_todo; t += 32) {
                const int j = todo[t];
                const double r = ((lo + (double)crow[j]) - ui) - v[j];
                double dj = dist[j];
                if (r < dj) { pred[j] = i; dist[j] = r; dj = r; }
                Cand c;
                c.d = dj; c.t = t; c.a = row4col[j] < 0;
                best = cand_combine(best, c);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                Cand c;
                c.d = __shfl_xor_sync(FULL, best.d, o);
                c.t = __shfl_xor_sync(FULL, best.t, o);
                c.a = __shfl_xor_sync(FULL, best.a, o);
                best = cand_combine(best, c);
            }
            lo = best.d;
            if (best.t < 0 || lo == INFINITY) { sink = -2; break; }  // infeasible (cannot happen for finite costs)
            const int j = todo[best.t];
            const int r4 = row4col[j];
            __syncwarp();
            if (lane == 0) { in_sc[j] = 1; todo[best.t] = todo[n_todo - 1]; }
            --n_todo;
            if (r4 < 0) sink = j; else i = r4;
            __syncwarp();
        }
        if (sink < 0) { if (lane == 0) atomicOr(p.status, 4); return; }
        // dual updates
        for (int r = lane; r < nr; r += 32) {
            if (r == cur) u[r] += lo;
            else if (in_sr[r]) u[r] += lo - dist[col4row[r]];
        }
        for (int j = lane; j < nc; j += 32)
            if (in_sc[j]) v[j] -= lo - dist[j];
        __syncwarp();
        if (lane == 0) {  // augment along the alternating path
            int j = sink;
            for (;;) {
                const int r = pred[j];
                row4col[j] = r;
                const int prev = col4row[r];
                col4row[r] = j;
                j = prev;
                if (r == cur) break;
            }
        }
        __syncwarp();
    }
    // dense assignment: query -> target
    for (int r = lane; r < nr; r += 32) {
        const int c = col4row[r];
        if (transposed) mt[c] = r; else mt[r] = c;
    }
    __syncwarp();
    if (p.pred_idx) {  // reference layout: prediction indices ascending (matcher.py:86-87)
        int base = 0;
        for (int q0 = 0; q0 < M; q0 += 32) {
            const int q = q0 + lane;
            const int t = q < M ? mt[q] : -1;
            const unsigned m = __ballot_sync(FULL, t >= 0);
            if (t >= 0) {
                const int k = base + __popc(m & ((1u << lane) - 1u));
                p.pred_idx[(long long)lb * p.KMAX + k] = q;
                p.tgt_idx[(long long)lb * p.KMAX + k] = t;
            }
            base += __popc(m);
        }
    }
}

static size_t match_fixed_smem(int M, int C, int NT) {
    const int ncmax = M > NT ? M : NT, nrmax = M < NT ? M : NT;
    size_t s = sizeof(double) * (nrmax + 2 * ncmax) + sizeof(int) * (3 * ncmax + nrmax) + (nrmax + ncmax) + 16;
    s += sizeof(float) * (size_t)M * C;
    return s;
}


// ------------------------------------------------------------------------------------------------ losses
struct LossParams {
    const float* logits;
    const float* boxes;
    const int32_t* match_tgt;
    const float* tgt_boxes;
    const int32_t* tgt_labels;
    const int32_t* tgt_count;
    const float* num_boxes;  // device scalar, already all-reduced / world and clamped to >= 1
    float* dlogits;          // (L,B,M,C) or NULL
    float* dboxes;           // (L,B,M,4) or NULL
    float* partial;          // (L,B,8): ce, l1, giou, card_pred, n_matched, n_correct
    int L, B, M, C, NT;
    float w_ce, w_bbox, w_giou, alpha;
};

__device__ __forceinline__ float softplusf_(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }

// d max(a,b)/da with torch's tie rule (half each)
__device__ __forceinline__ float gmax_a(float a, float b) { return a > b ? 1.f : (a == b ? 0.5f : 0.f); }
__device__ __forceinline__ float gmin_a(float a, float b) { return a < b ? 1.f : (a == b ? 0.5f : 0.f); }

__global__ void __launch_bounds__(MT_THREADS) set_loss_kernel(const LossParams p) {
    const int lb = blockIdx.x;
    const int b = lb % p.B;
    const int M = p.M, C = p.C;
    const float nb = *p.num_boxes;
    const float inv_nb = 1.f / nb;
    float s_ce = 0.f, s_l1 = 0.f, s_gi = 0.f, s_card = 0.f, s_nm = 0.f, s_ok = 0.f;
    for (int q = threadIdx.x; q < M; q += MT_THREADS) {
        const long long row = (long long)lb * M + q;
        const int t = p.match_tgt[row];
        int lab = -1;
        if (t >= 0) {
            lab = p.tgt_labels[(long long)b * p.NT + t];
            lab = lab < 0 ? 0 : (lab >= C ? C - 1 : lab);
        }
        // focal loss (segmentation.py:198-223), one-hot target from the assignment (set_criterion.py:44-55)
        float mx = -INFINITY;
        int arg = 0;
        for (int c = 0; c < C; ++c) {
            const float x = p.logits[row * C + c];
            if (x > mx) { mx = x; arg = c; }
            const float pr = sigmoidf_(x);
            float loss, g;
            if (c == lab) {
                const float om = 1.f - pr;
                const float nlogp = softplusf_(-x);
                loss = p.alpha * nlogp * om * om;
                g = p.alpha * om * om * (-2.f * pr * nlogp - om);
            } else {
                const float nlog1p = softplusf_(x);
                loss = (1.f - p.alpha) * nlog1p * pr * pr;
                g = (1.f - p.alpha) * pr * pr * (2.f * (1.f - pr) * nlog1p + pr);
            }
            s_ce += loss;
            if (p.dlogits) p.dlogits[row * C + c] = g * (p.w_ce * inv_nb);
        }
        if (mx > 0.5f) s_card += 1.f;  // set_criterion.py:85-88 (raw logits against 0.5, as written)
        float gb[4] = {0.f, 0.f, 0.f, 0.f};
        if (t >= 0) {
            s_nm += 1.f;
            if (arg == lab) s_ok += 1.f;
            const float* pb = p.boxes + row * 4;
            const float* tb = p.tgt_boxes + ((long long)b * p.NT + t) * 4;
            const float qcx = pb[0], qcy = pb[1], qw = pb[2], qh = pb[3];
            const float tcx = tb[0], tcy = tb[1], tw = tb[2], th = tb[3];
            s_l1 += fabsf(qcx - tcx) + fabsf(qcy - tcy) + fabsf(qw - tw) + fabsf(qh - th);
            const float sl = p.w_bbox * inv_nb;
            gb[0] = sl * ((qcx > tcx) - (qcx < tcx));
            gb[1] = sl * ((qcy > tcy) - (qcy < tcy));
            gb[2] = sl * ((qw > tw) - (qw < tw));
            gb[3] = sl * ((qh > th) - (qh < th));
            // GIoU (box_ops.py:32-69) forward + analytic backward
            const float x0 = qcx - 0.5f * qw, y0 = qcy - 0.5f * qh, x1 = qcx + 0.5f * qw, y1 = qcy + 0.5f * qh;
            const float u0 = tcx - 0.5f * tw, v0 = tcy - 0.5f * th, u1 = tcx + 0.5f * tw, v1 = tcy + 0.5f * th;
            const float ap = (x1 - x0) * (y1 - y0), at = (u1 - u0) * (v1 - v0);
            const float ix0 = fmaxf(x0, u0), ix1 = fminf(x1, u1), iy0 = fmaxf(y0, v0), iy1 = fminf(y1, v1);
            const float iwr = ix1 - ix0, ihr = iy1 - iy0;
            const float iw = fmaxf(iwr, 0.f), ih = fmaxf(ihr, 0.f);
            const float inter = iw * ih;
            const float uni = ap + at - inter;
            const float iou = inter / uni;
            const float cx0 = fminf(x0, u0), cx1 = fmaxf(x1, u1), cy0 = fminf(y0, v0), cy1 = fmaxf(y1, v1);
            const float cwr = cx1 - cx0, chr_ = cy1 - cy0;
            const float cw = fmaxf(cwr, 0.f), ch = fmaxf(chr_, 0.f);
            const float ac = cw * ch;
            const float giou = iou - (ac - uni) / ac;
            s_gi += 1.f - giou;
            const float g = -(p.w_giou * inv_nb);  // d total / d giou
            float d_uni = -inter / (uni * uni) * g + g / ac;
            float d_inter = g / uni - d_uni;
            const float d_ac = -uni / (ac * ac) * g;
            const float d_ap = d_uni;
            const float d_iw = d_inter * ih * (iwr >= 0.f), d_ih = d_inter * iw * (ihr >= 0.f);
            const float d_cw = d_ac * ch * (cwr >= 0.f), d_ch = d_ac * cw * (chr_ >= 0.f);
            float dx0 = -d_iw * gmax_a(x0, u0) - d_cw * gmin_a(x0, u0) - d_ap * (y1 - y0);
            float dx1 = d_iw * gmin_a(x1, u1) + d_cw * gmax_a(x1, u1) + d_ap * (y1 - y0);
            float dy0 = -d_ih * gmax_a(y0, v0) - d_ch * gmin_a(y0, v0) - d_ap * (x1 - x0);
            float dy1 = d_ih * gmin_a(y1, v1) + d_ch * gmax_a(y1, v1) + d_ap * (x1 - x0);
            gb[0] += dx0 + dx1;
            gb[1] += dy0 + dy1;
            gb[2] += 0.5f * (dx1 - dx0);
            gb[3] += 0.5f * (dy1 - dy0);
        }
        if (p.dboxes) {
            float4 o = make_float4(gb[0], gb[1], gb[2], gb[3]);
            reinterpret_cast<float4*>(p.dboxes)[row] = o;
        }
    }
    // block reduction (warp shuffles, then 4 warps through shared memory) — deterministic order
    __shared__ float red[MT_THREADS / 32][6];
    float vals[6] = {s_ce, s_l1, s_gi, s_card, s_nm, s_ok};
#pragma unroll
    for (int k = 0; k < 6; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) vals[k] += __shfl_xor_sync(0xffffffffu, vals[k], o);
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 6; ++k) red[threadIdx.x >> 5][k] = vals[k];
    __syncthreads();
    if (threadIdx.x < 6) {
        float s = 0.f;
        for (int w = 0; w < MT_THREADS / 32; ++w) s += red[w][threadIdx.x];
        p.partial[(long long)lb * 8 + threadIdx.x] = s;
    }
}

// out[0] = total loss; out[1..5] = labels, box_l1, box_giou, cardinality, class_error (st_detr.py:181-187);
// out[8 + 3*l + {0,1,2}] = unweighted loss_ce / loss_bbox / loss_giou of level l.
__global__ void set_loss_finalize_kernel(const float* __restrict__ partial, const int32_t* __restrict__ tgt_count,
                                         const float* __restrict__ num_boxes, float* __restrict__ out, int L, int B,
                                         int main_level, float w_ce, float w_bbox, float w_giou) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const float inv_nb = 1.f / *num_boxes;
    float total = 0.f;
    for (int l = 0; l < L; ++l) {
        float ce = 0.f, l1 = 0.f, gi = 0.f;
        for (int b = 0; b < B; ++b) {
            const float* pp = partial + ((long long)l * B + b) * 8;
            ce += pp[0]; l1 += pp[1]; gi += pp[2];
        }
        ce *= inv_nb; l1 *= inv_nb; gi *= inv_nb;
        out[8 + 3 * l] = ce; out[8 + 3 * l + 1] = l1; out[8 + 3 * l + 2] = gi;
        total += w_ce * ce + w_bbox * l1 + w_giou * gi;
        if (l == main_level) {
            float card = 0.f, nm = 0.f, ok = 0.f;
            for (int b = 0; b < B; ++b) {
                const float* pp = partial + ((long long)l * B + b) * 8;
                card += fabsf(pp[3] - (float)tgt_count[b]);
                nm += pp[4]; ok += pp[5];
            }
            out[1] = w_ce * ce; out[2] = w_bbox * l1; out[3] = w_giou * gi;
            out[4] = card / (float)B;
            out[5] = nm > 0.f ? 100.f - ok * (100.f / nm) : 100.f;
        }
    }
    out[0] = total;
}

}  // namespace fod

using namespace fod;

extern "C" int fod_targets_compact(const float* boxes_xyxy, const int64_t* classes, const int64_t* active,
                                   float* out_boxes, int32_t* out_labels, int32_t* out_count, float* total_count,
                                   int32_t B, int32_t slots, float inv_w, float inv_h, void* stream) {
    if (!boxes_xyxy || !classes || !active || !out_boxes || !out_labels || !out_count || B <= 0 || slots <= 0)
        return fod_fail(FOD_E_ARG, "targets_compact: null/empty");
    targets_compact_kernel<<<B, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        boxes_xyxy, classes, active, out_boxes, out_labels, out_count, total_count, slots, inv_w, inv_h);
    return fod_check_launch("targets_compact_kernel");
}

extern "C" int fod_match(const fod_match_args* a, void* stream) {
    if (!a || !a->logits || !a->boxes || !a->tgt_boxes || !a->tgt_labels || !a->tgt_count || !a->match_tgt ||
        !a->status)
        return fod_fail(FOD_E_ARG, "match: null pointer");
    if (a->L <= 0 || a->B <= 0 || a->M <= 0 || a->C <= 0 || a->NT <= 0 || a->M > 1024 || a->NT > 1024)
        return fod_fail(FOD_E_ARG, "match: bad shape");
    if ((a->pred_idx == nullptr) != (a->tgt_idx == nullptr)) return fod_fail(FOD_E_ARG, "match: pred_idx/tgt_idx");
    MatchParams p;
    p.logits = a->logits; p.boxes = a->boxes; p.tgt_boxes = a->tgt_boxes; p.tgt_labels = a->tgt_labels;
    p.tgt_count = a->tgt_count; p.match_tgt = a->match_tgt; p.pred_idx = a->pred_idx; p.tgt_idx = a->tgt_idx;
    p.status = a->status; p.cost_out = a->cost_out; p.cost_ws = a->cost_ws;
    p.L = a->L; p.B = a->B; p.M = a->M; p.C = a->C; p.NT = a->NT; p.KMAX = a->M < a->NT ? a->M : a->NT;
    p.w_class = a->w_class; p.w_bbox = a->w_bbox; p.w_giou = a->w_giou; p.alpha = a->alpha;
    const size_t fixed = match_fixed_smem(a->M, a->C, a->NT);
    const size_t budget = 200 * 1024;
    size_t cost_bytes = (size_t)a->M * a->NT * 4;
    if (fixed + cost_bytes > budget) {
        if (!a->cost_ws) return fod_fail(FOD_E_ARG, "match: cost block exceeds shared memory and no cost_ws given");
        cost_bytes = 0;
    }
    p.smem_cost_floats = (int)(cost_bytes / 4);
    const size_t smem = fixed + cost_bytes;
    static size_t attr = 0;
    if (smem > attr) {
        cudaFuncSetAttribute(match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr = smem;
    }
    match_kernel<<<a->L * a->B, MT_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(p);
    return fod_check_launch("match_kernel");
}

extern "C" int fod_set_loss(const fod_set_loss_args* a, void* stream) {
    if (!a || !a->logits || !a->boxes || !a->match_tgt || !a->tgt_boxes || !a->tgt_labels || !a->tgt_count ||
        !a->num_boxes || !a->partial || !a->out)
        return fod_fail(FOD_E_ARG, "set_loss: null pointer");
    if (a->L <= 0 || a->B <= 0 || a->M <= 0 || a->C <= 0 || a->NT <= 0 || a->L > 16)
        return fod_fail(FOD_E_ARG, "set_loss: bad shape");
    if (a->dboxes && (reinterpret_cast<uintptr_t>(a->dboxes) & 15)) return fod_fail(FOD_E_ARG, "set_loss: dboxes align");
    LossParams p;
    p.logits = a->logits; p.boxes = a->boxes; p.match_tgt = a->match_tgt; p.tgt_boxes = a->tgt_boxes;
    p.tgt_labels = a->tgt_labels; p.tgt_count = a->tgt_count; p.num_boxes = a->num_boxes;
    p.dlogits = a->dlogits; p.dboxes = a->dboxes; p.partial = a->partial;
    p.L = a->L; p.B = a->B; p.M = a->M; p.C = a->C; p.NT = a->NT;
    p.w_ce = a->w_ce; p.w_bbox = a->w_bbox; p.w_giou = a->w_giou; p.alpha = a->alpha;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    set_loss_kernel<<<a->L * a->B, MT_THREADS, 0, st>>>(p);
    int rc = fod_check_launch("set_loss_kernel");
    if (rc) return rc;
    set_loss_finalize_kernel<<<1, 32, 0, st>>>(a->partial, a->tgt_count, a->num_boxes, a->out, a->L, a->B,
                                               a->main_level, a->w_ce, a->w_bbox, a->w_giou);
    return fod_check_launch("set_loss_finalize_kernel");
}

// ---------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of fod_set_loss: the (L,B,8) partial-sum workspace lives in a file-scope device array instead of
// caller memory, for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).
// Every GPU that loads the library holds the array (128 KB) until exit.  In host code `g_set_loss_partial` is only a
// stand-in for the symbol, not a device address: the address is looked up once per device with that device current.
// All calls share the array and it keeps the last call's contents; the region a call uses is cleared on the caller's
// stream first.  Calls that overlap on the GPU (different streams, no ordering) race on it.
static __device__ __align__(256) float g_set_loss_partial[FOD_SET_LOSS_STATIC_MAX_LB * 8];

static int set_loss_static_scratch(float** out) {
    constexpr int kMaxDev = 64;
    static float* ptr[kMaxDev] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) return fod_fail(FOD_E_LAUNCH, "set_loss_static: cudaGetDevice failed");
    float* p = __atomic_load_n(&ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_set_loss_partial);
        if (e != cudaSuccess || !p) {
            (void)cudaGetLastError();
            return fod_fail(FOD_E_LAUNCH, "set_loss_static: cudaGetSymbolAddress(g_set_loss_partial) failed");
        }
        __atomic_store_n(&ptr[dev], p, __ATOMIC_RELEASE);  // the same value from every thread that gets here
    }
    *out = p;
    return FOD_OK;
}

extern "C" int fod_set_loss_static(const fod_set_loss_args* a, void* stream) {
    if (!a) return fod_fail(FOD_E_ARG, "set_loss_static: null args");
    if (a->L <= 0 || a->B <= 0 || (long long)a->L * a->B > FOD_SET_LOSS_STATIC_MAX_LB)
        return fod_fail(FOD_E_ARG, "set_loss_static: L*B exceeds FOD_SET_LOSS_STATIC_MAX_LB (4096); use fod_set_loss with a caller workspace");
    float* scratch = nullptr;
    int rc = set_loss_static_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (cudaMemsetAsync(scratch, 0, sizeof(float) * 8 * (size_t)a->L * a->B, st) != cudaSuccess) {
        (void)cudaGetLastError();
        return fod_fail(FOD_E_LAUNCH, "set_loss_static: clearing the scratch failed");
    }
    fod_set_loss_args b = *a;
    b.partial = scratch;  // a->partial is ignored
    return fod_set_loss(&b, stream);
}
