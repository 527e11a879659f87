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
} fod_linear_args;
int fod_linear_fwd(const fod_linear_args* a, void* stream);

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
} fod_attn_args;
int fod_attn_fwd(const fod_attn_args* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif
