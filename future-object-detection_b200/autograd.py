"""torch.autograd.Function wrappers: each block's forward AND backward run on the C-ABI kernels (ops.py).

Conventions: activations are 2-D (rows, channels) bf16, contiguous; `w` is the fp32 weight (a Parameter, or a
cat/pad of Parameters — it only routes the gradient), `wq` its bf16 copy that the kernels read.  Weight gradients
are produced in fp32 by the split-K MN-major GEMM (dW = dYᵀ·X, atomically reduced), activation gradients in bf16.
Train-mode dropout (ops.drop_site) is applied inside the GEMM epilogues and on the attention weights; the masks are
regenerated in backward from the site's (p, call) pair (include/fod_b200.h fod_dropout) — nothing is stored.  Parity with
the reference is defined with dropout disabled and statistical with it on.
"""
import torch
from torch.autograd import Function

from . import ops

BF16 = torch.bfloat16
_NUM_SMS = 148


def _bf(t):
    return t.to(BF16).contiguous() if t.dtype != BF16 else t.contiguous()


def weight_grad(dy, x):
    """dW (N, K) fp32 = dyᵀ (N, rows) · x (rows, K), contraction over rows on the MN-major tcgen05 GEMM; the library
    picks the split-K factor (plain stores when it does not split, zero-fill + vector reductions when it does)."""
    n, k = dy.shape[1], x.shape[1]
    return ops.gemm(dy, x, a_mn=True, b_mn=True, out_f32=True, split_k=-1)


# ---- gradients written straight into the flat gradient arena (optim.FlatTrainState) -------------------------------
# A Parameter that owns an arena slot carries `_fod_reducer`.  Its gradient kernels (split-K dW GEMM, column sums,
# LayerNorm backward) accumulate into `p.grad` in place and the Function returns None for it: no temporary, no zero
# fill, no AccumulateGrad add kernel.  `_fod_uses` counts the forward uses of the step so that the reducer's bucket
# countdown fires after the LAST contribution, as autograd's own accumulation would.
def _arena(p):
    return p is not None and getattr(p, "_fod_reducer", None) is not None and p.requires_grad


def _direct(p):
    """For module code (outside Function.forward, where grad mode is the caller's)."""
    return _arena(p) and torch.is_grad_enabled()


def _use(*ps):
    """Inside Function.forward of a call that will be differentiated: per-tensor flags; a use is counted for every
    arena-backed parameter."""
    flags = [_arena(p) for p in ps]
    for p, d in zip(ps, flags):
        if d:
            p._fod_uses = getattr(p, "_fod_uses", 0) + 1
    return flags


def _use_all(*ps):
    """All-or-nothing (block functions): direct mode only if EVERY parameter of the block owns an arena slot."""
    if all(_arena(p) for p in ps):
        _use(*ps)
        return True
    return False


def _done(p):
    p._fod_uses -= 1
    if p._fod_uses == 0:
        p._fod_reducer.mark_ready(p)


class FusedParams:
    """Several arena-backed parameters that optim.FlatTrainState laid out back to back (a module announces them through
    `fused_param_groups()`), seen as ONE parameter with the rows concatenated: `data`, `grad` and the bf16 operand copy are
    strided views over the arenas.  A fused projection (decoder q/k/v content halves, transformer.py:122-181) then
    needs no torch.cat of weights / biases per step, and its dW GEMM and bias column sum accumulate straight into the
    members' gradient slots instead of into a temporary that autograd splits and adds slice by slice."""
    requires_grad = True

    def __init__(self, members):
        p0 = members[0]
        shape = (sum(p.shape[0] for p in members),) + tuple(p0.shape[1:])
        stride = torch.empty(shape, device="meta").stride()
        self._fod_members = tuple(members)
        self._fod_reducer = p0._fod_reducer
        self._fod_uses = 0
        self.shape = shape
        self.data = torch.as_strided(p0.data, shape, stride)
        self.grad = torch.as_strided(p0.grad, shape, stride)
        self._fod_shadow = torch.as_strided(p0._fod_shadow, shape, stride)
        self.key = FusedParams.key_of(members)
        reg = getattr(self._fod_reducer, "fused", None)
        if reg is not None:
            reg[self.key] = self

    @staticmethod
    def key_of(members):
        return tuple((p.data_ptr(), p.grad.data_ptr() if p.grad is not None else 0) for p in members)

    @staticmethod
    def build(members, cached=None):
        """FusedParams over `members` if every one owns an arena slot, gradients are being recorded and the slots are
        adjacent in this order; else None (the caller falls back to torch.cat + autograd routing)."""
        if not torch.is_grad_enabled() or not all(_arena(p) and p.grad is not None for p in members):
            return None
        if cached is not None and cached.key == FusedParams.key_of(members):
            return cached
        for a, b in zip(members, members[1:]):
            if (b.data_ptr() != a.data_ptr() + 4 * a.numel() or b.grad.data_ptr() != a.grad.data_ptr() + 4 * a.numel()
                    or b._fod_shadow.data_ptr() != a._fod_shadow.data_ptr() + 2 * a.numel() or b.shape[1:] != a.shape[1:]):
                return None
        return FusedParams(members)


class _OffCriticalPath:
    """Weight / bias gradients are leaves of the backward graph: nothing downstream of backward waits for them except
    the gradient all-reduce and the optimizer.  They are enqueued on the reducer's weight-gradient stream, forked from
    the backward stream at the point where their operands exist, so that these small grids (a few CTAs each) fill SMs
    the activation-gradient chain leaves idle.  Under CUDA-graph capture the fork becomes a parallel branch.

    The operands the side stream reads are kept referenced until the backward stream has joined it again
    (reducer.finish): (1) the caching allocator cannot recycle them early, and (2) autograd's InputBuffer cannot
    accumulate another gradient INTO them in place (it only does so when it holds the last reference) — several of
    these operands are also returned as activation gradients (LayerNorm / ReLU-mask outputs)."""

    def __init__(self, reducer, *operands):
        self.ws = getattr(reducer, "wgrad_stream", None)
        self.reducer = reducer
        self.operands = operands

    def __enter__(self):
        if self.ws is not None:
            self.ws.wait_stream(torch.cuda.current_stream())
            self.ctx = torch.cuda.stream(self.ws)
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ws is not None:
            self.ctx.__exit__(*exc)
            self.reducer.keepalive.append(self.operands)


def weight_grad_into(p, dy, x, rows=None):
    g = p.grad.view(-1, x.shape[1])  # (N, K) view of the slot (a 1x1 conv weight is stored (N, K, 1, 1))
    if rows is not None:
        g = g[rows[0]:rows[1]]
    with _OffCriticalPath(p._fod_reducer, dy, x):
        ops.gemm(dy, x, a_mn=True, b_mn=True, accum_into=g)
    _done(p)


def bias_grad_into(p, dy):
    with _OffCriticalPath(p._fod_reducer, dy):
        ops.colsum(dy, out=p.grad.view(1, -1))
    _done(p)


def input_grad(dy, wq, **kw):  # kw: relu_mask / mask_scale / residual / colsum_out
    """dX (rows, K) bf16 = dy (rows, N) · W (N, K), W read in its nn.Linear layout (MN-major B operand)."""
    return ops.gemm(dy, wq, b_mn=True, **kw)


class LinearFn(Function):
    """y = [relu](x·Wᵀ + b + residual[row % mod]) — nn.Linear (+ReLU) call sites of transformer.py / paper.py."""

    @staticmethod
    def forward(ctx, x, w, b, wq, residual, res_row_mod, relu, out_f32, w_rows=None):
        """w_rows=(r0, r1): `wq` is rows [r0, r1) of the arena-backed parameter `w`; its gradient accumulates into the
        same rows of w.grad (direct mode only — otherwise pass the slice itself and let autograd route it)."""
        y = ops.gemm(x, wq, bias=b.data if isinstance(b, FusedParams) else b, residual=residual, res_row_mod=res_row_mod,
                     relu=relu, out_f32=out_f32)
        ctx.w_rows = w_rows
        ctx.save_for_backward(x, wq, y if relu else None)
        ctx.cfg = (res_row_mod, relu, b is not None, residual is not None)
        ctx.direct = _use(w, b) if any(ctx.needs_input_grad) else [False, False]
        ctx.params = (w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, wq, y = ctx.saved_tensors
        mod, relu, has_b, has_res = ctx.cfg
        dy = _bf(dy)
        if relu:
            dy = dy * (y > 0)
        dx = input_grad(dy, wq) if ctx.needs_input_grad[0] else None
        (w_direct, b_direct), (w, b) = ctx.direct, ctx.params
        dw = db = None
        if w_direct:
            weight_grad_into(w, dy, x, ctx.w_rows)
        elif ctx.w_rows is not None:
            raise RuntimeError("w_rows needs an arena-backed parameter")
        elif ctx.needs_input_grad[1]:
            dw = weight_grad(dy, x)
        if b_direct:
            bias_grad_into(b, dy)
        elif has_b and ctx.needs_input_grad[2]:
            db = ops.colsum(dy).view(-1)
        dres = None
        if has_res and ctx.needs_input_grad[4]:
            dres = dy if mod == 0 else ops.batch_sum(dy.view(-1, mod, dy.shape[1]))
        return dx, dw, db, None, dres, None, None, None, None


def linear(x, w, b, wq, residual=None, res_row_mod=0, relu=False, out_f32=False, w_rows=None):
    return LinearFn.apply(x, w, b, wq, residual, res_row_mod, relu, out_f32, w_rows)


class LinearLNFn(Function):
    """y = LayerNorm(res + x·Wᵀ + b): attention out_proj + residual + norm (transformer.py:417, 272, 286)."""

    @staticmethod
    def forward(ctx, x, res, w, b, wq, gamma, beta, eps, need_grad, drop=None):
        """drop: dropout on the sub-block output x·Wᵀ + b before the residual (transformer.py:271,285,417: dropout_sa /
        dropout_ia / attn_dropout)."""
        if need_grad:
            y, pre, mean, rstd = ops.gemm(x, wq, bias=b, residual=res, ln=(gamma, beta, eps), save_ln=True, drop=drop)
            ctx.save_for_backward(x, wq, pre, mean, rstd, gamma)
            ctx.direct = _use_all(w, b, gamma, beta)
            ctx.params = (w, b, gamma, beta)
            ctx.drop = drop
        else:
            y = ops.gemm(x, wq, bias=b, residual=res, ln=(gamma, beta, eps), drop=drop)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, wq, pre, mean, rstd, gamma = ctx.saved_tensors
        drop = ctx.drop
        if ctx.direct:  # all four parameters own arena slots: every gradient kernel accumulates in place
            w, b, g, bt = ctx.params
            dz = ops.ln_bwd(_bf(dy), pre, mean, rstd, gamma, g.grad, bt.grad, dzsum=b.grad, drop=drop)
            dz, du = dz if drop is not None else (dz, dz)  # gradients of the residual and of the (dropped-out) linear output
            dx = input_grad(du, wq)
            weight_grad_into(w, du, x)
            for q in (b, g, bt):
                _done(q)
            return dx, dz, None, None, None, None, None, None, None, None
        acc = torch.zeros((3, gamma.numel()), dtype=torch.float32, device=gamma.device)  # dgamma, dbeta, db
        dz = ops.ln_bwd(_bf(dy), pre, mean, rstd, gamma, acc[0], acc[1], dzsum=acc[2], drop=drop)
        dz, du = dz if drop is not None else (dz, dz)
        dx = input_grad(du, wq)
        dw = weight_grad(du, x)
        return dx, dz, dw, acc[2], None, acc[0], acc[1], None, None, None


def linear_ln(x, res, w, b, wq, gamma, beta, eps=1e-5, drop=None):
    return LinearLNFn.apply(x, res, w, b, wq, gamma, beta, eps, torch.is_grad_enabled(), drop)


class FFNLNFn(Function):
    """y = LayerNorm(x + W2·relu(W1·x + b1) + b2) — transformer.py:407-419 (encoder), :228-235,309-311 (decoder),
    :103-108,116-118 (egodeep MLP).  Backward: LN bwd -> dH GEMM with fused ReLU mask -> dX GEMM with fused residual."""

    @staticmethod
    def forward(ctx, x, w1, b1, w1q, w2, b2, w2q, gamma, beta, eps, need_grad, drop_h=None, drop_o=None):
        """drop_h: dropout on the hidden after the ReLU, drop_o: on the FFN output before the residual
        (transformer.py:407-412 mlp = Linear, ReLU, Dropout, Linear, Dropout; :228-235,310 decoder)."""
        h = ops.gemm(x, w1q, bias=b1, relu=True, drop=drop_h)  # saved AFTER dropout: its zeros cover both masks
        if need_grad:
            y, pre, mean, rstd = ops.gemm(h, w2q, bias=b2, residual=x, ln=(gamma, beta, eps), save_ln=True, drop=drop_o)
            ctx.save_for_backward(x, h, w1q, w2q, pre, mean, rstd, gamma)
            ctx.direct = _use_all(w1, b1, w2, b2, gamma, beta)
            ctx.params = (w1, b1, w2, b2, gamma, beta)
            ctx.drops = (drop_h, drop_o)
        else:
            y = ops.gemm(h, w2q, bias=b2, residual=x, ln=(gamma, beta, eps), drop=drop_o)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, h, w1q, w2q, pre, mean, rstd, gamma = ctx.saved_tensors
        drop_h, drop_o = ctx.drops
        hs = 1.0 / (1.0 - drop_h[0]) if drop_h is not None else 1.0  # kept hidden elements carry 1/(1-p)
        if ctx.direct:
            w1, b1, w2, b2, g, bt = ctx.params
            dz = ops.ln_bwd(_bf(dy), pre, mean, rstd, gamma, g.grad, bt.grad, dzsum=b2.grad, drop=drop_o)
            dz, du = dz if drop_o is not None else (dz, dz)
            dh = input_grad(du, w2q, relu_mask=h, colsum_out=b1.grad, mask_scale=hs)
            weight_grad_into(w2, du, h)
            dx = input_grad(dh, w1q, residual=dz)
            weight_grad_into(w1, dh, x)
            for q in (b1, b2, g, bt):
                _done(q)
            return dx, None, None, None, None, None, None, None, None, None, None, None, None
        d = gamma.numel()
        acc = torch.zeros((3 * d + h.shape[1],), dtype=torch.float32, device=gamma.device)  # dgamma, dbeta, db2, db1
        dgamma, dbeta, db2, db1 = acc[:d], acc[d:2 * d], acc[2 * d:3 * d], acc[3 * d:]
        dz = ops.ln_bwd(_bf(dy), pre, mean, rstd, gamma, dgamma, dbeta, dzsum=db2, drop=drop_o)
        dz, du = dz if drop_o is not None else (dz, dz)
        dh = input_grad(du, w2q, relu_mask=h, colsum_out=db1, mask_scale=hs)  # ReLU mask and the bias-1 gradient ride in the epilogue
        dw2 = weight_grad(du, h)
        dx = input_grad(dh, w1q, residual=dz)
        dw1 = weight_grad(dh, x)
        return dx, dw1, db1, None, dw2, db2, None, dgamma, dbeta, None, None, None, None


def ffn_ln(x, w1, b1, w1q, w2, b2, w2q, gamma, beta, eps=1e-5, drop_h=None, drop_o=None):
    return FFNLNFn.apply(x, w1, b1, w1q, w2, b2, w2q, gamma, beta, eps, torch.is_grad_enabled(), drop_h, drop_o)


class AddLNFn(Function):
    """y = LayerNorm(x + bcast[row // rows_per_group]) — norm_eda (transformer.py:486) and decoder.norm (:386)."""

    @staticmethod
    def forward(ctx, x, bcast, rows_per_group, gamma, beta, eps, need_grad):
        if need_grad:
            y, mean, rstd = ops.add_ln_fwd(x, gamma, beta, eps, bcast, rows_per_group, save=True)
            ctx.save_for_backward(x, bcast, mean, rstd, gamma)
            ctx.rpg = rows_per_group
            ctx.direct = _use_all(gamma, beta)
            ctx.params = (gamma, beta)
        else:
            y = ops.add_ln_fwd(x, gamma, beta, eps, bcast, rows_per_group)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, bcast, mean, rstd, gamma = ctx.saved_tensors
        if ctx.direct:
            g, bt = ctx.params
            dgamma, dbeta = g.grad, bt.grad
        else:
            dgamma = torch.zeros_like(gamma)
            dbeta = torch.zeros_like(gamma)
        dz = ops.ln_bwd(_bf(dy), x, mean, rstd, gamma, dgamma, dbeta, bcast=bcast, rows_per_group=ctx.rpg)
        db = ops.colsum(dz, ctx.rpg).to(BF16) if bcast is not None else None
        if ctx.direct:
            _done(g)
            _done(bt)
            return dz, db, None, None, None, None, None
        return dz, db, None, dgamma, dbeta, None, None


class AddLNDropFn(Function):
    """y = LayerNorm(x + dropout(x2)) — norm1(out + attn_dropout(out)) of EgodeepAttention (transformer.py:117, x2 is x)
    and norm_eda(features + dropout_eda(new)) of the encoder layer (:485-486).  Mask regenerated in backward."""

    @staticmethod
    def forward(ctx, x, x2, gamma, beta, eps, need_grad, drop):
        if need_grad:
            y, pre, mean, rstd = ops.add_ln_fwd_drop(x, x2, gamma, beta, eps, drop, save=True)
            ctx.save_for_backward(pre, mean, rstd, gamma)
            ctx.direct = _use_all(gamma, beta)
            ctx.params = (gamma, beta)
            ctx.drop = drop
        else:
            y = ops.add_ln_fwd_drop(x, x2, gamma, beta, eps, drop)
        return y

    @staticmethod
    def backward(ctx, dy):
        pre, mean, rstd, gamma = ctx.saved_tensors
        if ctx.direct:
            g, bt = ctx.params
            dgamma, dbeta = g.grad, bt.grad
        else:
            dgamma = torch.zeros_like(gamma)
            dbeta = torch.zeros_like(gamma)
        dz, dzd = ops.ln_bwd(_bf(dy), pre, mean, rstd, gamma, dgamma, dbeta, drop=ctx.drop)
        if ctx.direct:
            _done(g)
            _done(bt)
            return dz, dzd, None, None, None, None, None
        return dz, dzd, dgamma, dbeta, None, None, None


def add_ln_drop(x, x2, gamma, beta, eps=1e-5, drop=None):
    """drop must be a dropout site (ops.drop_site); without one use add_ln(x + x2)."""
    assert drop is not None
    return AddLNDropFn.apply(x, x2, gamma, beta, eps, torch.is_grad_enabled(), drop)


def add_ln(x, gamma, beta, bcast=None, rows_per_group=0, eps=1e-5):
    return AddLNFn.apply(x, bcast, rows_per_group, gamma, beta, eps, torch.is_grad_enabled())


class SelfAttnFn(Function):
    """Multi-head self-attention core on a fused projection qkv (B, L, 768) = [q | k | v], head dim 32:
    torch.nn.MultiheadAttention core (transformer.py:404) and SlotToSlotAttention's (transformer.py:79-81)."""

    @staticmethod
    def forward(ctx, qkv, heads, scale, drop=None):
        c = heads * 32
        out, lse = ops.attn_fwd(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], scale, heads, drop=drop)
        ctx.save_for_backward(qkv, out, lse)
        ctx.cfg = (heads, scale, drop)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        heads, scale, drop = ctx.cfg
        c = heads * 32
        dq, _, _, _, _ = ops.attn_bwd(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], out, lse, _bf(dout), scale,
                                      heads, fused="qkv", drop=drop)
        return dq._base if dq._base is not None else dq, None, None, None


def self_attention(qkv, heads=8, scale=32 ** -0.5, drop=None):
    """drop: dropout on the softmax weights (torch.nn.MultiheadAttention(dropout=...), attention.py:370)."""
    return SelfAttnFn.apply(qkv, heads, scale, drop)


class PlainAttnFn(Function):
    """Single-part attention on separate dense q (B,Lq,C), k, v (B,Lk,C) — EgodeepAttention with several IMU keys."""

    @staticmethod
    def forward(ctx, q, k, v, heads, scale, drop=None):
        out, lse = ops.attn_fwd(q, k, v, scale, heads, drop=drop)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (heads, scale, drop)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        heads, scale, drop = ctx.cfg
        dq, dk, dv, _, _ = ops.attn_bwd(q, k, v, out, lse, _bf(dout), scale, heads, drop=drop)
        return dq, dk, dv, None, None, None


class CrossAttnFn(Function):
    """Conditional cross-attention core (transformer.py:148-176 + attention.py:274-372): per head
    q = [qc ‖ qs], k = [kc ‖ kp] realised as two accumulating MMAs; kv (B, S, 512) = [kc | v] fused projection;
    kp (S, 256) batch-invariant; qs (B, M, 256) or batch-invariant (M, 256)."""

    @staticmethod
    def forward(ctx, qc, qs, kv, kp, heads, scale, drop=None):
        c = heads * 32
        out, lse = ops.attn_fwd(qc, kv[..., :c], kv[..., c:], scale, heads, q2=qs, k2=kp, batch=kv.shape[0], drop=drop)
        ctx.save_for_backward(qc, qs, kv, kp, out, lse)
        ctx.cfg = (heads, scale, drop)
        return out

    @staticmethod
    def backward(ctx, dout):
        qc, qs, kv, kp, out, lse = ctx.saved_tensors
        heads, scale, drop = ctx.cfg
        c = heads * 32
        dq1, dk1, _, dq2, dk2 = ops.attn_bwd(qc, kv[..., :c], kv[..., c:], out, lse, _bf(dout), scale, heads, q2=qs,
                                             k2=kp, fused="kv", drop=drop)
        dkv = dk1._base
        dqs = dq2 if qs.dim() == 3 else ops.batch_sum(dq2)
        dkp = ops.batch_sum(dk2) if kp.dim() == 2 else dk2
        return dq1, dqs, dkv, dkp, None, None, None


def cross_attention(qc, qs, kv, kp, heads=8, scale=64 ** -0.5, drop=None):
    return CrossAttnFn.apply(qc, qs, kv, kp, heads, scale, drop)


class SetLossFn(Function):
    """Matching + losses (set_criterion.py:172-217): forward runs fod_match and fod_set_loss (which also produces the
    gradients); backward scales them by the incoming gradient of the total loss."""

    @staticmethod
    def forward(ctx, logits, boxes, tgt_boxes, tgt_labels, tgt_count, num_boxes, weights, status):
        w_ce, w_bbox, w_giou, c_class, c_bbox, c_giou, alpha = weights
        m = ops.match(logits.detach(), boxes.detach(), tgt_boxes, tgt_labels, tgt_count, c_class, c_bbox, c_giou,
                      alpha, want_lists=False, status=status)
        out, dl, db = ops.set_loss(logits, boxes, m["match_tgt"], tgt_boxes, tgt_labels, tgt_count, num_boxes, w_ce,
                                   w_bbox, w_giou, alpha, want_grads=True)
        ctx.save_for_backward(dl, db)
        ctx.mark_non_differentiable(m["match_tgt"])
        return out, m["match_tgt"]

    @staticmethod
    def backward(ctx, dout, _):
        dl, db = ctx.saved_tensors
        g = dout[0]
        return dl * g, db * g, None, None, None, None, None, None
