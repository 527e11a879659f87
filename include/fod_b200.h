/* fod_b200 C ABI — B200 (sm_100a) kernels for the hot path of atonderski/future-object-detection.
 *
 * Every entry point takes raw device pointers, explicit sizes and a CUDA stream (as void*), returns an int
 * status (0 = ok, see FOD_E_*), never allocates, never synchronises the device.  The reference is pure
 * Python/PyTorch, so there is no FFI in it to mirror; each function below names the reference code
 * (file:line under /root/reference) whose arithmetic it replaces.  The Python binding a maintainer would
 * add is a ctypes stub (see INTEGRATION.md, and future-object-detection_b200/lib.py which is that stub).
 *
 * Tensor conventions: activations are bf16 (uint16 storage), row-major, "tokens x channels" with the batch
 * dimension outermost (batch-first; the reference is sequence-first — layouts are converted at the module
 * boundary).  Statistics, biases, LayerNorm parameters, losses, boxes and logits are fp32.  Indices int64.
 */
#ifndef FOD_B200_H
#define FOD_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FOD_OK 0
#define FOD_E_ARG 1      /* bad shape / alignment / null pointer */
#define FOD_E_TMAP 2     /* TMA descriptor creation failed */
#define FOD_E_LAUNCH 3   /* kernel launch failed (cudaGetLastError) */
#define FOD_E_NUMERIC 4  /* matcher: NaN / -inf in cost matrix (scipy raises ValueError) */

int fod_version(void);                  /* ABI version, bumped on signature changes */
const char* fod_last_error(void);       /* thread-local message for the last non-zero status */

/* ------------------------------------------------------------------------------------------------
 * Train-mode dropout (p = 0.1 in the reference: on the attention weights, ConditionalDETR/models/attention.py:370 and
 * torch.nn.MultiheadAttention; after every sub-block and inside the FFNs, future_od/models/transformer.py:201,207,234,
 * 271,285,310,405-412).  Masks are never stored: forward and backward regenerate them with Philox4x32-7 from
 * rng_state = device uint64[2] {seed, step}, the call-site id `call` and the element index.  `step` is advanced on the
 * device once per training step (fod_rng_advance), so a captured CUDA graph draws fresh masks on every replay.
 * RNG streams cannot match torch's: parity with the reference is statistical (keep rate, scale), exact at p = 0.
 */
typedef struct fod_dropout {
    const void* rng_state;   /* NULL or p == 0: no dropout */
    float p;                 /* drop probability in [0, 1) */
    int32_t call;            /* call-site id, unique per dropout site within a step */
} fod_dropout;
int fod_rng_advance(void* rng_state, void* stream);   /* rng_state[1] += 1 (one thread) */

/* ------------------------------------------------------------------------------------------------
 * Linear + fused epilogue on tcgen05 tensor cores:  Y = epi(X · Wᵀ)
 *   epi: (+bias) (+residual[row % res_row_mod]) (ReLU) (LayerNorm over the N=256 row) -> bf16 and/or fp32
 * Replaces: nn.Linear / F.relu / residual add / nn.LayerNorm chains in
 *   future_od/models/transformer.py:407-419 (encoder FFN + norms), :228-235,309-311 (decoder FFN),
 *   :18-32 (MLP), :75-79,148-174 (attention input projections), ConditionalDETR/models/attention.py:373
 *   (out_proj), torch.nn.MultiheadAttention in-proj (transformer.py:404).
 * X (M,K) bf16 row stride ldx; W (N,K) bf16 contiguous (nn.Linear layout); K % 8 == 0.
 * LayerNorm requires N == 256.  pre_ln / ln_mean / ln_rstd are optional training-time outputs.
 */
typedef struct fod_linear_args {
    const void* x;
    const void* w;
    const float* bias;      /* (N) or NULL */
    const void* residual;   /* bf16 (rows, N) or NULL */
    const float* ln_gamma;  /* (N) or NULL -> no LayerNorm */
    const float* ln_beta;
    void* out_bf16;         /* (M,N) or NULL */
    float* out_f32;         /* (M,N) or NULL */
    void* pre_ln_bf16;      /* (M,N) or NULL: value before LayerNorm */
    float* ln_mean;         /* (M) or NULL */
    float* ln_rstd;         /* (M) or NULL */
    int64_t ldx, ld_res, ld_out;
    int32_t M, N, K;
    int32_t res_row_mod;    /* 0: residual row = row; >0: row % res_row_mod (batch-invariant positional term) */
    int32_t relu;
    float ln_eps;
    /* --- backward / general-GEMM extensions (zero = forward defaults) --- */
    const void* relu_mask;  /* bf16 (M,N), ld_out: output zeroed where mask <= 0 (ReLU backward, transformer.py:408) */
    const float* col_scale; /* fp32 (res_row_mod or M, N): y = (acc + bias) * col_scale  (query_scale ⊙ sine, :366) */
    int64_t ldw;            /* row stride of w; 0 = contiguous */
    int32_t a_mn;           /* 1: x is stored (K, M) row-major (contraction outermost) — dW = dYᵀ·X */
    int32_t b_mn;           /* 1: w is stored (K, N) row-major — dX = dY·W with W in its nn.Linear layout */
    int32_t split_k;        /* >1: split the contraction over grid.z (needs atomic_out); <0: the library picks the
                               split and, without atomic_out, zero-fills out_f32 itself when it splits */
    int32_t atomic_out;     /* 1: out_f32 += result (vector reductions into a caller-initialised buffer) */
    float* colsum_out;      /* (N) fp32 or NULL: += column sums of the stored bf16 output — the bias gradient of the
                               layer that produced the GEMM's input gradient (transformer.py:407-410 backward) */
    fod_dropout drop;       /* dropout on (acc + bias [ReLU]) — x + dropout(sublayer(x)) with the LayerNorm epilogue (the
                               residual is never dropped), dropout(relu(.)) for an FFN hidden; N % 8 == 0, no residual
                               outside the LayerNorm epilogue */
    float mask_scale;       /* with relu_mask: kept elements are multiplied by it (0 = 1): 1/(1-p) when the saved hidden
                               is a dropped-out ReLU output, whose zeros cover both masks */
    int32_t pad1;
} fod_linear_args;
int fod_linear_fwd(const fod_linear_args* a, void* stream);  /* also the backward GEMMs via a_mn / b_mn */

/* ------------------------------------------------------------------------------------------------
 * Flash attention forward (TMA + tcgen05/TMEM, online softmax), multi-head, no mask:
 *   S = scale * (Q1·K1ᵀ [+ Q2·K2ᵀ]);  P = softmax(S);  O = P·V
 * The optional second (Q2,K2) pair is the ConditionalDETR "concat" form [q_content ‖ q_sine]·[k_content ‖ k_pos]
 * done by accumulation instead of torch.cat.
 * Replaces: torch.nn.MultiheadAttention core (future_od/models/transformer.py:404,417) and
 *   ConditionalDETR/models/attention.py:274-372 (scaling, bmm, softmax, bmm) for
 *   SlotToSlotAttention (transformer.py:61-82) and SlotToImageAttention (transformer.py:122-181).
 * Layouts: q* (B, Lq, H*32) bf16, k* (B, Lk, H*32), v (B, Lk, H*32), out (B, Lq, H*32) bf16,
 *   lse (B, H, Lq) fp32 (natural-log sum-exp of the scaled scores), head dim fixed at 32 per part.
 */
typedef struct fod_attn_args {
    const void* q1;
    const void* k1;
    const void* q2;  /* NULL for single-part attention */
    const void* k2;
    const void* v;
    void* out;
    float* lse;
    int64_t ldq1, ldk1, ldq2, ldk2, ldv, ldo;          /* row strides in elements */
    int64_t bsq1, bsk1, bsq2, bsk2, bsv, bso;          /* batch strides in elements (0 = broadcast) */
    int32_t B, H, Lq, Lk;
    float scale;
    int32_t pad0;
    fod_dropout drop;  /* dropout on the softmax weights P (out = dropout(P)·V; lse is that of the undropped P); the
                          backward call must carry the same descriptor */
    const void* keep_bits;  /* optional: the keep decisions of `drop` as written by fod_attn_keep_bits (NULL: the kernels
                               draw them themselves; the bits are a cache of the same function, never a different mask) */
    int64_t keep_pitch;     /* bytes per (b, head, query) row of keep_bits */
} fod_attn_args;
int fod_attn_fwd(const fod_attn_args* a, void* stream);
/* Keep decisions of the dropout on the softmax weights (ConditionalDETR/models/attention.py:370; torch.nn.
 * MultiheadAttention in future_od/models/transformer.py:404) for one attention call, drawn ONCE for the forward and the
 * backward kernel: out[(b*H + h)*Lq + q][k / 8] bit (k % 8) = 1 if P[b, h, q, k] is kept under (rng_state, call, p) —
 * exactly the decision the kernels draw on their own when keep_bits is NULL.  rows = B*H*Lq; pitch bytes per row, a
 * multiple of 16 with pitch >= 16*ceil(Lk/128); out 16-byte aligned. */
int fod_attn_keep_bits(const void* rng_state, int32_t call, float p, int64_t rows, int32_t Lk, void* out, int64_t pitch,
                       void* stream);


/* Flash attention backward (same layouts as fod_attn_fwd; recomputes S and dP on the tensor cores):
 *   delta = rowsum(dO ⊙ O) (written to `delta`, (B,H,Lq) fp32 workspace), then
 *   dQ_p = scale·(P ⊙ (dP − delta))·K_p,  dK_p = scale·(P ⊙ (dP − delta))ᵀ·Q_p,  dV = Pᵀ·dO.
 * Gradients are dense bf16: dq1/dq2 (B,Lq,H*32), dk1/dk2/dv (B,Lk,H*32) — a batch-broadcast K part (bsk2 == 0)
 * still gets a per-batch dk2 that the caller sums.  Backward of ConditionalDETR/models/attention.py:350-372 and of
 * torch.nn.MultiheadAttention's core (future_od/models/transformer.py:404,417).
 */
typedef struct fod_attn_bwd_args {
    fod_attn_args fwd;   /* the forward call's arguments: q/k/v, out (= O), lse, strides, scale */
    const void* dout;    /* (B,Lq,H*32) bf16 */
    int64_t lddo, bsdo;
    float* delta;        /* (B,H,Lq) fp32 workspace/output */
    void *dq1, *dq2, *dk1, *dk2, *dv;
    int64_t ld_dq;       /* row stride of dq1 (0 = H*32); batch stride = Lq*ld_dq.  dq2 is always dense */
    int64_t ld_dk;       /* row stride of dk1 and dv (0 = H*32); batch stride = Lk*ld_dk.  dk2 is always dense */
    int32_t mode;        /* 0: the library picks; 1: two kernels (dQ per query tile, dK/dV per key tile: S and dP are
                            recomputed twice, any length); 2: ONE fused kernel, one recompute, dQ of every query tile
                            resident in TMEM — Lq <= 768 (one part) / Lq <= 256 (two parts) per CTA; longer one-part
                            sequences run it in query groups of 768 and need `ws` */
    int32_t pad0;
    float* ws;           /* optional fp32 workspace of >= 2*B*H*Lk*32 floats, 16-byte aligned (may be NULL): with it, one-part
                            calls with Lq > 768 take the fused kernel in query groups (dK / dV partial sums are reduced
                            in `ws` with fp32 vector atomics — 2 addends at Lq <= 1536, so bit-reproducible there — and
                            then written as bf16); without it they take the two-kernel path */
    int64_t ws_floats;   /* capacity of ws in floats */
} fod_attn_bwd_args;
int fod_attn_bwd(const fod_attn_bwd_args* a, void* stream);

/* Row kernels over 256-channel rows (HBM-bound):
 *   fod_add_ln_fwd: y = LayerNorm(x + bcast[row / rows_per_group]) — norm_eda over the broadcast egodeep vector
 *     (future_od/models/transformer.py:479-486) and the decoder's per-level norm (:386-387; bcast = NULL).
 *   fod_ln_bwd: dz, dgamma += , dbeta += for z = pre (+ bcast[row / rows_per_group]); dy2 is an optional second
 *     addend of the incoming gradient.  Backward of every nn.LayerNorm on the path.
 *   fod_colsum: out[g][c] += sum_{rows of group g} x[row][c] (bias and broadcast-operand gradients).
 */
int fod_add_ln_fwd(const void* x, const void* bcast, int32_t rows_per_group, const float* gamma, const float* beta,
                   void* y, float* mean_out, float* rstd_out, int64_t M, float eps, void* stream);
int fod_ln_bwd(const void* dy, const void* dy2, const void* pre, const void* bcast, int32_t rows_per_group,
               const float* mean, const float* rstd, const float* gamma, void* dz, float* dgamma, float* dbeta,
               float* dzsum /* (256) or NULL: += column sums of dz = bias gradient of the linear feeding the norm */,
               int64_t M, void* stream);
int fod_colsum(const void* x, int64_t ldx, float* out, int64_t M, int32_t N, int32_t rows_per_group, void* stream);
/* y = LayerNorm(x + dropout(x2) [+ bcast[row / rows_per_group]]): the residual + dropout + norm of blocks whose
 * sub-layer output is not the epilogue of a GEMM — EgodeepAttention's norm1(out + attn_dropout(out)) (x2 == x) and
 * the encoder layer's norm_eda(features + dropout_eda(new)) (future_od/models/transformer.py:117,485-486).  x2 may alias
 * x; pre_out (optional) receives the pre-LayerNorm sum that fod_ln_bwd_drop takes as `pre`; the mask is the one
 * fod_ln_bwd_drop regenerates (element index row*256 + col of `drop`). */
int fod_add_ln_fwd_drop(const void* x, const void* x2, const void* bcast, int32_t rows_per_group, const float* gamma,
                        const float* beta, void* y, void* pre_out, float* mean_out, float* rstd_out, int64_t M, float eps,
                        const fod_dropout* drop, void* stream);
/* fod_ln_bwd for y = LayerNorm(res + dropout(u)): dz is the gradient of res, dz_drop (optional) = dz masked and scaled
 * by the SAME dropout mask the forward GEMM epilogue drew (`drop`: element index row*256 + col) = the gradient of u;
 * dzsum then sums dz_drop (the bias sits inside the dropout). */
int fod_ln_bwd_drop(const void* dy, const void* dy2, const void* pre, const void* bcast, int32_t rows_per_group,
                    const float* mean, const float* rstd, const float* gamma, void* dz, void* dz_drop, float* dgamma,
                    float* dbeta, float* dzsum, int64_t M, const fod_dropout* drop, void* stream);
/* 3x3 / stride 2 / pad 1 max pooling, NHWC bf16, C % 8 == 0: the ResNet stem pool in front of the frozen layer1
 * (ConditionalDETR/models/backbone.py:88-96 -> torchvision resnet `maxpool`); forward only (paper.py:102-109 freezes
 * everything before layer2).  y is (N, (H-1)/2+1, (W-1)/2+1, C). */
int fod_maxpool3x3s2_nhwc(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, void* stream);
/* out[i] = sum over g < G of x[g*n + i] (bf16 in / out, fp32 accumulation; n % 8 == 0): gradient of a batch-invariant
 * operand — the positional projections every sequence shares (future_od/models/transformer.py:459-460,160-166,
 * paper.py:370) entered the GEMMs as row-modular residuals and the attention calls as broadcast parts, so their
 * gradient is the sum over the batch of the per-sequence gradient. */
int fod_batch_sum(const void* x, void* out, int64_t G, int64_t n, void* stream);
/* Input frames for the backbone: video (B, L, 3, H, W) fp32 planar, of which frames [0, T) of every clip are used (the
 * model drops the last, "future" frame: future_od/models/paper.py:463-468) -> out (B*T, H, W, 3) bf16 = a channels-last
 * (B*T, 3, H, W) tensor, what the frozen stem consumes.  stride_b / stride_t: element strides between clips / frames;
 * planes are dense (H*W), W % 8 == 0.  One pass instead of torch's slice copy + channels_last copy + cast. */
int fod_frames_to_nhwc_bf16(const float* video, int64_t stride_b, int64_t stride_t, int32_t B, int32_t T, int32_t H,
                            int32_t W, void* out, void* stream);
/* The same frames in the layout fod_stem_conv7x7s2_p4 reads: out (B*T, H, W + 8, 4) bf16 — pixel x at padded column
 * x + 3, the fourth channel and the 3 left / 5 right border columns zero (all written by this call).  W % 8 == 0. */
int fod_frames_to_nhwc4p_bf16(const float* video, int64_t stride_b, int64_t stride_t, int32_t B, int32_t T, int32_t H,
                              int32_t W, void* out, void* stream);

/* Fused ReLU backward with an optional gradient add for the backbone's bottleneck blocks: out = (a [+ b]) * (y > 0) on
 * flat bf16 arrays of n elements with identical layouts (n % 8 == 0, 16-byte aligned).  Replaces autograd's gradient
 * accumulation (block input used by conv1 and by the identity path, torchvision Bottleneck.forward as built by
 * future_od/models/paper.py:94-98) followed by threshold_backward with one pass. */
int fod_relu_bwd_add(const void* a, const void* b /* or NULL */, const void* y, void* out, int64_t n, void* stream);

/* ResNet stem: conv 7x7 / stride 2 / pad 3, 3 -> 64 channels, frozen BatchNorm folded, ReLU
 * (ConditionalDETR/models/backbone.py:88-96 -> torchvision resnet conv1+bn1+relu; frozen, forward only: paper.py:102-109).
 * x (N,H,W,3) bf16 NHWC, W % 8 == 0; wk (64,192) bf16 with wk[o][kh*24 + 1 + kw*3 + c] = W[o][c][kh][kw]*scale[o] and
 * zeros elsewhere; bias (64) fp32 = folded shift; y (N,(H-1)/2+1,(W-1)/2+1,64) bf16 NHWC.  Implicit GEMM on tcgen05. */
int fod_stem_conv7x7s2(const void* x, const void* wk, const float* bias, void* y, int32_t N, int32_t H, int32_t W,
                       void* stream);
/* The same convolution with the operand tile fetched by TMA instead of assembled by the threads (same reference lines).
 * x4p (N, H, W + 8, 4) bf16 as written by fod_frames_to_nhwc4p_bf16 (H even, W % 8 == 0); wk4 (64, 224) bf16 with
 * wk4[o][kh*32 + kw*4 + c] = W[o][c][kh][kw]*scale[o] for kw < 7, c < 3 and zeros elsewhere; bias (64) fp32;
 * y (N, H/2, W/2, 64) bf16 NHWC.  One overlapping-window tensor map yields the im2col rows of a kernel row directly. */
int fod_stem_conv7x7s2_p4(const void* x4p, const void* wk4, const float* bias, void* y, int32_t N, int32_t H, int32_t W,
                          void* stream);
/* The stem convolution AND the max-pool that follows it (torchvision resnet `maxpool`, kernel 3 / stride 2 / pad 1:
 * ConditionalDETR/models/backbone.py:88-96; frozen, forward only: paper.py:102-109) in one kernel: the 64-channel
 * convolution output stays on chip, y is the pooled (N, (H/2-1)/2+1, (W/2-1)/2+1, 64) bf16 NHWC tensor.  Operands as
 * for fod_stem_conv7x7s2_p4.  Equal to fod_maxpool3x3s2_nhwc(fod_stem_conv7x7s2_p4(x)) bit for bit. */
int fod_stem_conv_pool_p4(const void* x4p, const void* wk4, const float* bias, void* y, int32_t N, int32_t H, int32_t W,
                          void* stream);

/* Frozen-BatchNorm folding of the trainable backbone convolutions (ConditionalDETR/models/backbone.py:53-63 applies
 * y = conv(x, W)*s + t with frozen s, t; the cuDNN conv runs on W*s).  One descriptor per conv, the table lives in
 * DEVICE memory: out[o][kh][kw][i] = bf16(w[o][i][kh][kw] * scale[o])  (torch channels_last storage). */
typedef struct fod_fold_desc {
    const float* w;      /* (O, I, KH, KW) fp32, contiguous */
    const float* scale;  /* (O) */
    void* out;           /* bf16, O*KH*KW*I elements */
    int32_t O, I, HW, pad;
} fod_fold_desc;
/* ONE launch for `count` convs; max_elems = the largest O*I*HW in the table (grid sizing). */
int fod_fold_conv_weights(const fod_fold_desc* descs_device, int32_t count, int64_t max_elems, void* stream);
/* Backward of the fold: grad[o][i][hw] += float(dwq[o][hw][i]) * scale[o]; dwq is cuDNN's bf16 channels_last weight
 * gradient of the folded conv, grad the fp32 (O, I, KH, KW) gradient slot of the unfolded weight. */
int fod_conv_wgrad_unfold(const void* dwq, const float* scale, float* grad, int32_t O, int32_t I, int32_t HW,
                          void* stream);

/* ------------------------------------------------------------------------------------------------
 * AP bookkeeping of one forward pass: future_od/utils/od_map.py:214-287 prepare_od_map_stuffs,
 * called from future_od/models/st_detr.py:222-229 inside every forward.  One launch, grid (B, C):
 *   top-K detections per class by score; IoU+1e-7 smoothing against the N annotation slots (od_map.py:48-71); the
 *   K-step greedy claim loop at T IoU thresholds (od_map.py:267-277); size buckets (od_map.py:163-181,199-211).
 * scores carry the generic "any class" column last (st_detr.py:200-205); class c < C-1 only claims annotations of
 * class c, the generic column claims any active annotation.  Slots with active != 1 never match, so the caller need
 * not drop unused slots (od_map.py:40-45) — outputs are identical.
 * Outputs: confs (T, C, B*K) fp32; is_positive (T, C, B*K) and size_cat (C, 4, B*K) as 1-byte bools;
 * num_annos (C, 4) int64, += (zero it first).  Limits: M, N <= 256, K <= min(64, M), T <= 10.
 */
typedef struct fod_ap_args {
    const float* pred_boxes;
    const float* scores;
    const float* anno_boxes;
    const void* anno_classes;   /* int64 (B, N) */
    const void* anno_active;    /* int64 (B, N) */
    const float* thresholds;    /* (T) */
    float* confs;
    void* is_positive;
    void* size_cat;
    void* num_annos;
    int32_t B, M, C, N, K, T;
    float size_lo, size_hi;     /* bucket delimiters in pixel^2: (1/24)(1/64)·H·W and (1/4)(1/12)·H·W */
} fod_ap_args;
int fod_ap_match(const fod_ap_args* a, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Flat optimizer step.  One contiguous arena of n fp32 elements holds every trainable parameter
 * (n % 4 == 0, 16-byte slots); grads / exp_avg / exp_avg_sq / the bf16 shadow mirror its layout.  One call =
 *   total = ||grads||_2; coef = min(1, max_norm / (total + 1e-6))      torch.nn.utils.clip_grad_norm_, trainer.py:186
 *   t += 1; AdamW(lr, betas, eps, weight_decay) on coef * grads          torch.optim.AdamW.step, trainer.py:188 with the
 *                                                                         parameter groups of runs/_helper.py:84-107
 *   shadow_bf16 = bf16(params)                                           the operand copies the GEMM kernels read
 *   grads = 0 (zero_grad != 0) or coef * grads                           optimizer.zero_grad, trainer.py:171
 * in three launches (sum of squares -> scalars -> update), no host synchronisation.
 * state: 8 device floats {grad norm, clip coef, step t, 1 - beta1^t, 1 - beta2^t, ...}; state[2] must start at 0.
 * partial: fod_optim_workspace_floats() device floats.  Groups: elements [group_begin[i], group_begin[i+1]) use
 * group_lr[i] / group_wd[i].
 */
typedef struct fod_adamw_args {
    float* params;
    float* grads;
    float* exp_avg;
    float* exp_avg_sq;
    void* shadow_bf16;   /* (n) bf16 or NULL */
    float* state;
    float* partial;
    int64_t n;
    int64_t group_begin[4];
    float group_lr[4];
    float group_wd[4];
    int32_t num_groups;
    int32_t zero_grad;
    float beta1, beta2, eps, max_norm;
} fod_adamw_args;
int fod_flat_adamw_step(const fod_adamw_args* a, void* stream);
int fod_optim_workspace_floats(void);
int fod_cast_bf16(const float* src, void* dst, int64_t n, void* stream);  /* dst = bf16(src), n % 4 == 0 */

/* ------------------------------------------------------------------------------------------------
 * Set-prediction loss path (north-star piece 4).  Levels are the decoder levels (main output = level L-1).
 *
 * fod_targets_compact — replaces to_detr_targets (future_od/models/st_detr.py:237-263): dense 256-slot xyxy-pixel
 *   annotations -> per-sample compacted cxcywh boxes normalised by (W,H), labels and counts; no host sync.
 *   inv_w / inv_h are float32(1/W), float32(1/H) as the reference builds them.  total_count (optional, device
 *   float, must be zeroed by the caller) accumulates sum(Nt) for num_boxes (set_criterion.py:185-193).
 */
int fod_targets_compact(const float* boxes_xyxy /*(B,slots,4)*/, const int64_t* classes /*(B,slots)*/,
                        const int64_t* active /*(B,slots)*/, float* out_boxes /*(B,slots,4)*/,
                        int32_t* out_labels /*(B,slots)*/, int32_t* out_count /*(B)*/, float* total_count,
                        int32_t B, int32_t slots, float inv_w, float inv_h, void* stream);

/* fod_match — replaces HungarianMatcher.forward (ConditionalDETR/models/matcher.py:40-87), generalized_box_iou /
 *   box_cxcywh_to_xyxy (ConditionalDETR/util/box_ops.py:17-69) and scipy.optimize.linear_sum_assignment
 *   (matcher.py:86) for all L levels and B samples in one launch.  Only the per-sample diagonal blocks are built.
 *   match_tgt[l,b,q] = matched target index or -1.  pred_idx/tgt_idx (optional) are the reference's index lists
 *   (prediction indices ascending), padded with -1 to KMAX = min(M,NT).  *status is OR'd with
 *   1 (NaN/-inf cost: scipy's ValueError), 2 (degenerate box: the assert at box_ops.py:59-60), 4 (infeasible).
 */
typedef struct fod_match_args {
    const float* logits;       /* (L,B,M,C) */
    const float* boxes;        /* (L,B,M,4) cxcywh in [0,1] */
    const float* tgt_boxes;    /* (B,NT,4) compacted */
    const int32_t* tgt_labels; /* (B,NT) */
    const int32_t* tgt_count;  /* (B) */
    int32_t* match_tgt;        /* (L,B,M) out */
    int64_t* pred_idx;         /* (L,B,KMAX) out or NULL */
    int64_t* tgt_idx;          /* (L,B,KMAX) out or NULL */
    int32_t* status;           /* (1) device flag word, caller zeroes */
    float* cost_out;           /* (L,B,M,NT) optional copy of the fp32 cost blocks (tests) */
    float* cost_ws;            /* (L,B,M,NT) workspace, needed only when M*NT*4 + bookkeeping > 200 KB */
    int32_t L, B, M, C, NT;
    float w_class, w_bbox, w_giou, alpha;
} fod_match_args;
int fod_match(const fod_match_args* a, void* stream);

/* fod_set_loss — replaces SetCriterion.loss_labels/loss_boxes/loss_cardinality (future_od/models/set_criterion.py:
 *   36-115), sigmoid_focal_loss (ConditionalDETR/models/segmentation.py:198-223), accuracy (util/misc.py:439-455) and
 *   the weighting of st_detr.py:169-188, forward AND backward: dlogits/dboxes are d(total)/d(input).
 *   out[0] total; out[1..5] = labels, box_l1, box_giou, cardinality, class_error; out[8+3l..] per-level
 *   unweighted loss_ce, loss_bbox, loss_giou.  out needs 8+3L floats, partial L*B*8 floats.
 */
typedef struct fod_set_loss_args {
    const float* logits;
    const float* boxes;
    const int32_t* match_tgt;
    const float* tgt_boxes;
    const int32_t* tgt_labels;
    const int32_t* tgt_count;
    const float* num_boxes;    /* device scalar: clamp(sum Nt over ranks / world, 1) */
    float* dlogits;            /* (L,B,M,C) or NULL */
    float* dboxes;             /* (L,B,M,4) or NULL */
    float* partial;            /* (L,B,8) workspace */
    float* out;                /* (8+3L) */
    int32_t L, B, M, C, NT, main_level;
    float w_ce, w_bbox, w_giou, alpha;
} fod_set_loss_args;
int fod_set_loss(const fod_set_loss_args* a, void* stream);

/* fod_set_loss_static — no counterpart in the reference; added for callers that
 *   cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 *   cudaMalloc).  Same computation and outputs as fod_set_loss, but the (L,B,8) partial-sum workspace is a file-scope
 *   device array inside the library (128 KB per GPU that loads it, held until exit) and `a->partial` is ignored; the
 *   call allocates nothing.  Limit: L*B <= FOD_SET_LOSS_STATIC_MAX_LB, larger problems are rejected with FOD_E_ARG.
 *   All calls on a device share the array: it is cleared on `stream` at the start of every call, and two calls that
 *   may overlap on the GPU (different streams without ordering) race — serialise them or use fod_set_loss.
 */
#define FOD_SET_LOSS_STATIC_MAX_LB 4096
int fod_set_loss_static(const fod_set_loss_args* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif
