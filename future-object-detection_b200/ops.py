"""Raw operator layer: torch CUDA tensors -> C-ABI calls (include/fod_b200.h).  No autograd here (see autograd.py).

Every function enqueues on torch's current stream, allocates its outputs with torch (device memory plumbing only)
and raises `lib.FodError` on a non-zero status.  `LAUNCH_COUNTS` counts the kernels each call launches so
bench.py can report `gpu_launches`.
"""
import collections
import ctypes as C
import math

import os

import torch

from . import lib
from . import profile as prof

LAUNCH_COUNTS = collections.Counter()
BF16 = torch.bfloat16


# ---------------------------------------------------------------------------------------------- dropout RNG state
# One device-resident {seed, step} pair per device (include/fod_b200.h fod_dropout).  A dropout site is a (p, call)
# pair: `call` numbers the sites of one forward pass in execution order (the backward of a site reuses its number, saved
# in the autograd context), `step` is advanced on the device by rng_advance() once per training forward, so a captured
# CUDA graph — whose kernel arguments, `call` included, are frozen — draws fresh masks on every replay.
_RNG = {}
_CALL = [0]
_SNAP = [None]
_BITS_USED = []   # attention keep-bit keys of the current forward pass, in call order
_BITS_SIDE = {}   # device -> stream the keep bits of a pass are drawn on


def rng_state(device):
    key = str(device)
    if key not in _RNG:
        _RNG[key] = torch.tensor([0x5EED5EED, 0], dtype=torch.int64, device=device)
    return _RNG[key]


def set_dropout_seed(seed, device="cuda"):
    """Seed of the dropout masks (default 0x5EED5EED); also resets the step counter."""
    st = rng_state(torch.device(device) if not isinstance(device, torch.device) else device)
    st.copy_(torch.tensor([int(seed), 0], dtype=torch.int64))


def begin_forward(training=False):
    """Start of a model forward pass (FuturePredCore.forward): dropout call ids restart at 1, and the first dropout
    site of the pass advances the device-side step counter and takes a SNAPSHOT of {seed, step}, so every pass — and
    every replay of a captured graph — draws fresh masks.  The snapshot travels with each site's (p, call id) into the
    autograd context: backward regenerates its masks from it, so several forward passes may precede their backward
    passes (gradient accumulation) without mixing up masks."""
    _CALL[0] = 0
    _SNAP[0] = None
    # attention keep bits (see _attn_keep_bits): what the previous forward pass needed is drawn for THIS pass up front, on a
    # side stream, where it overlaps the backbone instead of sitting in front of every attention kernel
    plan = list(_BITS_USED)
    del _BITS_USED[:]
    if training and plan and _ATTN_BITS and _BITS_PREGEN and torch.cuda.is_available():
        dev = torch.device("cuda", torch.cuda.current_device())
        rng_advance(dev)                       # (what the first dropout site of the pass would have done)
        snap = _SNAP[0] = rng_state(dev).clone()
        side = _BITS_SIDE.get(str(dev))
        if side is None:
            side = _BITS_SIDE[str(dev)] = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for key in plan:
                _draw_keep_bits(snap, key, dev)
        ev = torch.cuda.Event()
        ev.record(side)
        snap._fod_bits_ready = ev


def end_forward():
    """End of a model forward pass: joins the keep-bit side stream if no attention call of the pass did (a pass that
    turned out to have no dropout site), so that no forked work is left dangling — a CUDA-graph capture requires it."""
    snap = _SNAP[0]
    ev = getattr(snap, "_fod_bits_ready", None) if snap is not None else None
    if ev is not None:
        torch.cuda.current_stream().wait_event(ev)
        snap._fod_bits_ready = None


def drop_site(p, training):
    """(p, call id, state snapshot) of the next dropout site of this forward pass, or None when dropout is off."""
    if not training or not p or p <= 0.0:
        return None
    if not 0.0 < p < 1.0:
        raise ValueError("dropout probability has to be in [0, 1), got %r" % (p,))
    if _SNAP[0] is None:
        dev = torch.device("cuda", torch.cuda.current_device())
        rng_advance(dev)
        _SNAP[0] = rng_state(dev).clone()
    _CALL[0] += 1
    return (float(p), _CALL[0], _SNAP[0])


def rng_advance(device):
    L = lib.load()
    lib.check(L.fod_rng_advance(_p(rng_state(device)), _stream()), "fod_rng_advance")
    LAUNCH_COUNTS["rng_advance_kernel"] += 1


def _fill_drop(d, drop, device):
    """drop = (p, call id[, state tensor]): without a state tensor the device's live {seed, step} is read."""
    if drop is not None:
        st = drop[2] if len(drop) > 2 and drop[2] is not None else rng_state(device)
        d.rng_state, d.p, d.call = _p(st), drop[0], drop[1]


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk_bf16(t, name):
    if t.dtype != BF16 or not t.is_cuda:
        raise lib.FodError(f"{name}: expected a CUDA bf16 tensor, got {t.dtype} on {t.device}")


def gemm(x, w, bias=None, residual=None, res_row_mod=0, relu=False, ln=None, out_f32=False, save_ln=False,
         relu_mask=None, col_scale=None, a_mn=False, b_mn=False, split_k=1, accum_into=None, M=None, N=None, K=None,
         colsum_out=None, out=None, drop=None, mask_scale=1.0):
    """Y = epi(op(x) · op(w)ᵀ) on the tcgen05 GEMM.  x (M,K) [or (K,M) if a_mn], w (N,K) [or (K,N) if b_mn].

    Returns y (bf16, or fp32 if out_f32); with ln=(gamma, beta, eps) and save_ln also (pre_ln, mean, rstd).
    accum_into: fp32 (M,N) tensor that receives vector reductions (weight gradients accumulated in place; the library
    picks the split-K factor).  split_k=-1 with out_f32: library-chosen split-K with overwrite semantics.
    colsum_out: fp32 (N,) that receives += column sums of the bf16 output (a fused bias gradient).
    drop=(p, call): dropout on (acc + bias [ReLU]) — before the residual under the LayerNorm epilogue; mask_scale: factor
    on the elements a relu_mask keeps (1/(1-p) when the mask is a dropped-out ReLU output)."""
    L = lib.load()
    _chk_bf16(x, "gemm x")
    _chk_bf16(w, "gemm w")
    assert x.dim() == 2 and w.dim() == 2 and x.stride(1) == 1 and w.stride(1) == 1
    if M is None:
        M, K = (x.shape[1], x.shape[0]) if a_mn else (x.shape[0], x.shape[1])
    if N is None:
        N = w.shape[1] if b_mn else w.shape[0]
    a = lib.LinearArgs()
    a.x, a.w = _p(x), _p(w)
    a.ldx, a.ldw = x.stride(0), w.stride(0)
    a.M, a.N, a.K = M, N, K
    a.a_mn, a.b_mn = int(a_mn), int(b_mn)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N
        a.bias = _p(bias)
    if residual is not None:
        _chk_bf16(residual, "gemm residual")
        assert residual.stride(-1) == 1
        a.residual, a.ld_res, a.res_row_mod = _p(residual), residual.stride(0), res_row_mod
    a.relu = int(relu)
    if relu_mask is not None:
        _chk_bf16(relu_mask, "gemm relu_mask")
        assert relu_mask.is_contiguous() and relu_mask.shape == (M, N)
        a.relu_mask = _p(relu_mask)
        a.mask_scale = mask_scale
    _fill_drop(a.drop, drop, x.device)
    if col_scale is not None:
        assert col_scale.dtype == torch.float32 and col_scale.is_contiguous()
        a.col_scale = _p(col_scale)
        a.res_row_mod = res_row_mod
    outs = []
    if accum_into is not None:
        assert accum_into.dtype == torch.float32 and accum_into.is_contiguous() and accum_into.shape == (M, N)
        a.out_f32, a.atomic_out, a.split_k = _p(accum_into), 1, (split_k if split_k > 1 else -1)
        y = accum_into
    elif out_f32:
        y = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=x.device)
        a.out_f32 = _p(y)
        if split_k < 0:
            a.split_k = -1
    else:
        y = torch.empty((M, N), dtype=BF16, device=x.device)
        a.out_bf16 = _p(y)
    a.ld_out = N
    if colsum_out is not None:
        assert colsum_out.dtype == torch.float32 and colsum_out.numel() == N and accum_into is None and not out_f32
        a.colsum_out = _p(colsum_out)
    outs.append(y)
    if ln is not None:
        gamma, beta, eps = ln
        a.ln_gamma, a.ln_beta, a.ln_eps = _p(gamma), _p(beta), eps
        if save_ln:
            pre = torch.empty((M, N), dtype=BF16, device=x.device)
            mean = torch.empty((M,), dtype=torch.float32, device=x.device)
            rstd = torch.empty((M,), dtype=torch.float32, device=x.device)
            a.pre_ln_bf16, a.ln_mean, a.ln_rstd = _p(pre), _p(mean), _p(rstd)
            outs += [pre, mean, rstd]
    _t = prof.begin()
    lib.check(L.fod_linear_fwd(C.byref(a), _stream()), "fod_linear_fwd")
    prof.end(_t, "%s%s_%dx%dx%d" % ("gemm_dw" if a_mn else ("gemm_dx" if b_mn else "gemm_fwd"), "_ln" if ln is not None else "",
                                    M, N, K), 2.0 * M * N * K,
             2.0 * (M * K + N * K) + M * N * (4 if (out_f32 or accum_into is not None) else 2),
             replay=(lambda: L.fod_linear_fwd(C.byref(a), _stream())) if _t is not None else None,
             keep=(x, w, bias, residual, relu_mask, col_scale, accum_into, colsum_out, ln, outs, drop))
    LAUNCH_COUNTS["gemm_bf16_kernel"] += 1
    return outs[0] if len(outs) == 1 else tuple(outs)


def _seq(t, B):
    """(ptr, row stride, batch stride) of a (B, L, C) or batch-invariant (L, C) bf16 tensor with unit channel stride."""
    _chk_bf16(t, "attention operand")
    assert t.stride(-1) == 1
    if t.dim() == 2:
        return _p(t), t.stride(0), 0
    assert t.dim() == 3 and t.shape[0] in (1, B)
    return _p(t), t.stride(1), (t.stride(0) if t.shape[0] == B and B > 1 else 0)


def _fill_attn(a, q1, k1, v, q2, k2, out, lse, scale, heads, drop=None):
    B, Lq = out.shape[0], out.shape[1]
    Lk = v.shape[-2]
    a.q1, a.ldq1, a.bsq1 = _seq(q1, B)
    a.k1, a.ldk1, a.bsk1 = _seq(k1, B)
    a.v, a.ldv, a.bsv = _seq(v, B)
    if q2 is not None:
        a.q2, a.ldq2, a.bsq2 = _seq(q2, B)
        a.k2, a.ldk2, a.bsk2 = _seq(k2, B)
    a.out, a.ldo, a.bso = _p(out), out.stride(1), out.stride(0)
    a.lse = _p(lse)
    a.B, a.H, a.Lq, a.Lk, a.scale = B, heads, Lq, Lk, scale
    _fill_drop(a.drop, drop, out.device)
    bits = _attn_keep_bits(drop, B, heads, Lq, Lk, out.device)
    if bits is not None:
        a.keep_bits, a.keep_pitch = _p(bits), bits.shape[1]


_ATTN_BITS = os.environ.get("FOD_ATTN_BITS", "1") != "0"  # 0: every attention kernel draws its dropout mask itself
_BITS_PREGEN = os.environ.get("FOD_ATTN_BITS_PREGEN", "1") != "0"  # 0: bits drawn at first use, on the compute stream


def _draw_keep_bits(st, key, device):
    """fod_attn_keep_bits for one attention call -> (rows, pitch) uint8, stored in the snapshot's cache."""
    call_id, p, B, heads, Lq, Lk = key
    cache = getattr(st, "_fod_keep_bits", None)
    if cache is None:
        cache = st._fod_keep_bits = {}
    L = lib.load()
    rows, pitch = B * heads * Lq, 16 * ((Lk + 127) // 128)
    bits = torch.empty((rows, pitch), dtype=torch.uint8, device=device)
    _t = prof.begin()
    call = lambda: L.fod_attn_keep_bits(_p(st), call_id, p, rows, Lk, _p(bits), pitch, _stream())
    lib.check(call(), "fod_attn_keep_bits")
    prof.end(_t, "attn_keep_bits_%dx%d" % (rows, Lk), 0.0, float(rows * pitch), replay=call if _t is not None else None,
             keep=(st, bits))
    LAUNCH_COUNTS["attn_keep_bits_kernel"] += 1
    cache[key] = bits
    return bits


def _attn_keep_bits(drop, B, heads, Lq, Lk, device):
    """The keep decisions of one attention call's dropout as a (B*heads*Lq, pitch) byte tensor (fod_attn_keep_bits),
    drawn once and shared by the forward and the backward kernel: cached on the forward's {seed, step} snapshot, which
    both carry in their drop tuple and which is never written again.  None without a snapshot (the live state moves).
    From the second training pass on the bits are already there: begin_forward() drew them on a side stream."""
    if not _ATTN_BITS or drop is None or len(drop) < 3 or drop[2] is None or not drop[0] > 0:
        return None
    st = drop[2]
    key = (drop[1], float(drop[0]), B, heads, Lq, Lk)
    if st is _SNAP[0] and key not in _BITS_USED:
        _BITS_USED.append(key)
    bits = getattr(st, "_fod_keep_bits", {}).get(key)
    if bits is None:
        return _draw_keep_bits(st, key, device)
    ev = getattr(st, "_fod_bits_ready", None)
    if ev is not None:  # first consumer of bits drawn on the side stream: join it (later consumers follow in stream order)
        torch.cuda.current_stream().wait_event(ev)
        st._fod_bits_ready = None
    return bits


def attn_fwd(q1, k1, v, scale, heads=8, q2=None, k2=None, batch=None, drop=None):
    """Flash attention forward.  Operands (B, L, heads*32) bf16 (row-strided views allowed) or (L, heads*32) when
    batch-invariant.  Returns out (B, Lq, heads*32) bf16 and lse (B, heads, Lq) fp32."""
    L = lib.load()
    B = batch if batch is not None else max(t.shape[0] for t in (q1, k1, v) if t.dim() == 3)
    Lq = q1.shape[-2]
    out = torch.empty((B, Lq, heads * 32), dtype=BF16, device=v.device)
    lse = torch.empty((B, heads, Lq), dtype=torch.float32, device=v.device)
    a = lib.AttnArgs()
    _fill_attn(a, q1, k1, v, q2, k2, out, lse, scale, heads, drop)
    _t = prof.begin()
    lib.check(L.fod_attn_fwd(C.byref(a), _stream()), "fod_attn_fwd")
    Lk, parts, D = v.shape[-2], (2 if q2 is not None else 1), heads * 32
    kb = B if (k2 is not None and k2.dim() == 3) else 1
    prof.end(_t, "attn_fwd_p%d_%dx%d" % (parts, Lq, Lk), (2.0 + 2.0 * parts) * B * Lq * Lk * D,
             2.0 * D * (B * Lq * (parts + 1) + B * Lk * 2 + (parts - 1) * kb * Lk),
             replay=(lambda: L.fod_attn_fwd(C.byref(a), _stream())) if _t is not None else None,
             keep=(q1, k1, v, q2, k2, out, lse, drop))
    LAUNCH_COUNTS["attn_fwd_kernel"] += 1
    return out, lse


def attn_bwd_is_fused(Lq, two_parts, mode=0):
    """Which kernel path fod_attn_bwd takes (mirrors csrc/attn_bwd_sm100.cu): the one-kernel backward needs the dQ
    accumulators of every 128-query tile resident in TMEM."""
    import os
    # one-part calls beyond 6 query tiles run the fused kernel in query groups (attn_bwd always hands it a workspace)
    fits = (Lq + 127) // 128 <= (2 if two_parts else 6) or not two_parts
    if mode:
        return mode == 2
    return fits and int(os.environ.get("FOD_ATTN_BWD_FUSED", "1")) != 0


def attn_bwd(q1, k1, v, out, lse, dout, scale, heads=8, q2=None, k2=None, fused=None, mode=0, drop=None):
    """Flash attention backward -> (dq1, dk1, dv, dq2, dk2), (B, L, heads*32) bf16 (dq2/dk2 None if unused).
    mode: 0 library picks, 1 two-kernel path, 2 fused single kernel (include/fod_b200.h).

    fused="qkv": dq1/dk1/dv are column slices of ONE (B, L, 3*heads*32) tensor (self-attention on a fused projection);
    fused="kv": dk1/dv are slices of one (B, Lk, 2*heads*32) tensor.  The kernels write the slices in place."""
    L = lib.load()
    B, Lq = out.shape[0], out.shape[1]
    Lk = v.shape[-2]
    dev = out.device
    g = lib.AttnBwdArgs()
    _fill_attn(g.fwd, q1, k1, v, q2, k2, out, lse, scale, heads, drop)
    _chk_bf16(dout, "attn dout")
    assert dout.shape == out.shape and dout.stride(-1) == 1
    g.dout, g.lddo, g.bsdo = _p(dout), dout.stride(1), dout.stride(0)
    delta = torch.empty((B, heads, Lq), dtype=torch.float32, device=dev)
    Cw = heads * 32
    if fused == "qkv":
        assert Lq == Lk
        whole = torch.empty((B, Lq, 3 * Cw), dtype=BF16, device=dev)
        dq1, dk1, dv = whole[..., :Cw], whole[..., Cw:2 * Cw], whole[..., 2 * Cw:]
        g.ld_dq = g.ld_dk = 3 * Cw
    elif fused == "kv":
        dq1 = torch.empty((B, Lq, Cw), dtype=BF16, device=dev)
        whole = torch.empty((B, Lk, 2 * Cw), dtype=BF16, device=dev)
        dk1, dv = whole[..., :Cw], whole[..., Cw:]
        g.ld_dk = 2 * Cw
    else:
        dq1 = torch.empty((B, Lq, Cw), dtype=BF16, device=dev)
        dk1 = torch.empty((B, Lk, Cw), dtype=BF16, device=dev)
        dv = torch.empty((B, Lk, Cw), dtype=BF16, device=dev)
    dq2 = dk2 = None
    g.delta, g.dq1, g.dk1, g.dv = _p(delta), _p(dq1), _p(dk1), _p(dv)
    if q2 is not None:
        dq2 = torch.empty((B, Lq, Cw), dtype=BF16, device=dev)
        dk2 = torch.empty((B, Lk, Cw), dtype=BF16, device=dev)
        g.dq2, g.dk2 = _p(dq2), _p(dk2)
    g.mode = mode
    ws = None
    if q2 is None and (Lq + 127) // 128 > 6 and mode != 1:  # query groups: fp32 dK / dV partial sums (include/fod_b200.h)
        ws = torch.empty((2 * B * heads * Lk * 32,), dtype=torch.float32, device=dev)
        g.ws, g.ws_floats = _p(ws), ws.numel()
    _t = prof.begin()
    lib.check(L.fod_attn_bwd(C.byref(g), _stream()), "fod_attn_bwd")
    parts = 2 if q2 is not None else 1
    prof.end(_t, "attn_bwd_p%d_%dx%d" % (parts, Lq, Lk), (4.0 + 6.0 * parts) * B * Lq * Lk * Cw,
             2.0 * Cw * (B * Lq * (2 * parts + 2) + B * Lk * (2 * parts + 2)),
             replay=(lambda: L.fod_attn_bwd(C.byref(g), _stream())) if _t is not None else None,
             keep=(q1, k1, v, q2, k2, out, lse, dout, delta, dq1, dk1, dv, dq2, dk2, drop, ws))
    if attn_bwd_is_fused(Lq, q2 is not None, mode):
        LAUNCH_COUNTS["attn_bwd_fused_kernel"] += 1
        if ws is not None:
            LAUNCH_COUNTS["attn_ws_zero_kernel"] += 1
            LAUNCH_COUNTS["attn_ws_to_bf16_kernel"] += 1
    else:
        LAUNCH_COUNTS["attn_delta_kernel"] += 1
        LAUNCH_COUNTS["attn_bwd_dq_kernel"] += 1
        LAUNCH_COUNTS["attn_bwd_dkdv_kernel"] += 1
    return dq1, dk1, dv, dq2, dk2


def add_ln_fwd(x, gamma, beta, eps=1e-5, bcast=None, rows_per_group=0, save=False):
    """y = LayerNorm(x + bcast[row // rows_per_group]); x (M,256) bf16.  save -> also (mean, rstd)."""
    L = lib.load()
    _chk_bf16(x, "add_ln x")
    assert x.is_contiguous() and x.shape[-1] == 256
    M = x.numel() // 256
    y = torch.empty_like(x)
    mean = rstd = None
    if save:
        mean = torch.empty((M,), dtype=torch.float32, device=x.device)
        rstd = torch.empty((M,), dtype=torch.float32, device=x.device)
    if bcast is not None:
        _chk_bf16(bcast, "add_ln bcast")
        assert bcast.is_contiguous() and bcast.numel() * rows_per_group == x.numel()
    _t = prof.begin()
    # the replay closure holds raw pointers only: a tensor captured here would keep this step's autograd graph alive
    ptrs = (_p(x), _p(bcast), rows_per_group, _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), M, eps)
    call = lambda: L.fod_add_ln_fwd(*ptrs, _stream())
    lib.check(call(), "fod_add_ln_fwd")
    prof.end(_t, "add_ln_fwd_%d" % M, 0.0, 1024.0 * M, replay=call if _t is not None else None,
             keep=(x, bcast, gamma, beta, y, mean, rstd))
    LAUNCH_COUNTS["add_ln_fwd_kernel"] += 1
    return (y, mean, rstd) if save else y


def add_ln_fwd_drop(x, x2, gamma, beta, eps=1e-5, drop=None, save=False):
    """y = LayerNorm(x + dropout(x2)); x, x2 (M,256) bf16 (x2 may be x).  save -> also (pre, mean, rstd) for ln_bwd."""
    L = lib.load()
    _chk_bf16(x, "add_ln_drop x")
    _chk_bf16(x2, "add_ln_drop x2")
    assert x.is_contiguous() and x2.is_contiguous() and x.shape == x2.shape and x.shape[-1] == 256
    M = x.numel() // 256
    y = torch.empty_like(x)
    pre = mean = rstd = None
    if save:
        pre = torch.empty_like(x)
        mean = torch.empty((M,), dtype=torch.float32, device=x.device)
        rstd = torch.empty((M,), dtype=torch.float32, device=x.device)
    dstruct = lib.Dropout()
    _fill_drop(dstruct, drop, x.device)
    _t = prof.begin()
    ptrs = (_p(x), _p(x2), None, 0, _p(gamma), _p(beta), _p(y), _p(pre), _p(mean), _p(rstd), M, eps)
    call = lambda: L.fod_add_ln_fwd_drop(*ptrs, C.byref(dstruct), _stream())
    lib.check(call(), "fod_add_ln_fwd_drop")
    prof.end(_t, "add_ln_fwd_drop_%d" % M, 0.0, (2048.0 if save else 1536.0) * M, replay=call if _t is not None else None,
             keep=(x, x2, gamma, beta, y, pre, mean, rstd, drop))
    LAUNCH_COUNTS["add_ln_fwd_kernel"] += 1
    return (y, pre, mean, rstd) if save else y


def ln_bwd(dy, pre, mean, rstd, gamma, dgamma, dbeta, dy2=None, bcast=None, rows_per_group=0, dzsum=None, drop=None):
    """LayerNorm backward: returns dz (bf16); accumulates into fp32 dgamma / dbeta (256,) and, if given, the column
    sums of dz into dzsum (256,) — the bias gradient of the linear layer that feeds the norm.
    drop=(p, call) of the forward GEMM epilogue (y = LN(res + dropout(u))): returns (dz, dz_drop) — the gradients of res
    and of u — and dzsum sums dz_drop."""
    L = lib.load()
    _chk_bf16(dy, "ln_bwd dy")
    _chk_bf16(pre, "ln_bwd pre")
    assert dy.is_contiguous() and pre.is_contiguous() and pre.shape[-1] == 256
    M = pre.numel() // 256
    dz = torch.empty_like(pre)
    dzd = torch.empty_like(pre) if drop is not None else None
    dstruct = lib.Dropout()
    _fill_drop(dstruct, drop, pre.device)
    _t = prof.begin()
    ptrs = (_p(dy), _p(dy2), _p(pre), _p(bcast), rows_per_group, _p(mean), _p(rstd), _p(gamma), _p(dz), _p(dzd), _p(dgamma),
            _p(dbeta), _p(dzsum), M)
    call = lambda: L.fod_ln_bwd_drop(*ptrs, C.byref(dstruct), _stream())
    lib.check(call(), "fod_ln_bwd")
    prof.end(_t, "ln_bwd_%d" % M, 0.0, 1536.0 * M, replay=call if _t is not None else None,
             keep=(dy, dy2, pre, bcast, mean, rstd, gamma, dz, dzd, dgamma, dbeta, dzsum, drop))
    LAUNCH_COUNTS["ln_bwd_kernel"] += 1
    return dz if drop is None else (dz, dzd)


def colsum(x, rows_per_group=0, out=None):
    """out[g, c] (+)= sum over rows of group g of x[row, c]; x (M,N) bf16, unit column stride."""
    L = lib.load()
    _chk_bf16(x, "colsum x")
    assert x.dim() == 2 and x.stride(1) == 1
    M, N = x.shape
    rpg = rows_per_group or M
    if out is None:
        out = torch.zeros((M // rpg, N), dtype=torch.float32, device=x.device)
    _t = prof.begin()
    ptrs = (_p(x), x.stride(0), _p(out), M, N, rpg)
    call = lambda: L.fod_colsum(*ptrs, _stream())
    lib.check(call(), "fod_colsum")
    prof.end(_t, "colsum_%dx%d" % (M, N), 0.0, 2.0 * M * N, replay=call if _t is not None else None, keep=(x, out))
    LAUNCH_COUNTS["colsum_kernel"] += 1
    return out


def maxpool3x3s2_nhwc(x):
    """nn.MaxPool2d(3, 2, 1) on a channels_last bf16 (N,C,H,W) tensor -> channels_last (N,C,OH,OW); forward only."""
    L = lib.load()
    _chk_bf16(x, "maxpool x")
    n, c, h, w = x.shape
    assert x.is_contiguous(memory_format=torch.channels_last) and c % 8 == 0
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, c, oh, ow), dtype=BF16, device=x.device, memory_format=torch.channels_last)
    _t = prof.begin()
    ptrs = (_p(x), _p(y), n, h, w, c)
    call = lambda: L.fod_maxpool3x3s2_nhwc(*ptrs, _stream())
    lib.check(call(), "fod_maxpool3x3s2_nhwc")
    prof.end(_t, "maxpool_%dx%dx%d" % (n, h, w), 0.0, 2.0 * (x.numel() + y.numel()), replay=call if _t is not None else None,
             keep=(x, y))
    LAUNCH_COUNTS["maxpool3x3s2_nhwc_kernel"] += 1
    return y


def batch_sum(x):
    """x (G, ...) bf16 contiguous -> sum over dim 0 (fp32 accumulation, bf16 result): fod_batch_sum."""
    L = lib.load()
    _chk_bf16(x, "batch_sum x")
    assert x.dim() >= 2
    G, n = x.shape[0], x[0].numel()
    if n % 8 or G == 1 or not x.is_contiguous() or x.data_ptr() % 16:
        return x.sum(0) if G > 1 else x[0]
    out = torch.empty(x.shape[1:], dtype=BF16, device=x.device)
    _t = prof.begin()
    ptrs = (_p(x), _p(out), G, n)
    call = lambda: L.fod_batch_sum(*ptrs, _stream())
    lib.check(call(), "fod_batch_sum")
    prof.end(_t, "batch_sum_%dx%d" % (G, n), 0.0, 2.0 * (G + 1) * n, replay=call if _t is not None else None, keep=(x, out))
    LAUNCH_COUNTS["batch_sum_kernel"] += 1
    return out


def frames_to_nhwc_bf16(video, frames=None):
    """video (B, L, 3, H, W) fp32 with dense (3, H, W) frames -> the first `frames` (default L) frames of every clip as a
    channels-last bf16 (B*frames, 3, H, W) tensor, in one pass (fod_frames_to_nhwc_bf16)."""
    L = lib.load()
    b, l, c, h, w = video.shape
    t = l if frames is None else frames
    assert video.dtype == torch.float32 and c == 3 and 0 < t <= l and video.stride()[2:] == (h * w, w, 1) and w % 8 == 0
    out = torch.empty((b * t, h, w, 3), dtype=BF16, device=video.device)
    _t = prof.begin()
    args = (_p(video), video.stride(0), video.stride(1), b, t, h, w, _p(out))
    call = lambda: L.fod_frames_to_nhwc_bf16(*args, _stream())
    lib.check(call(), "fod_frames_to_nhwc_bf16")
    prof.end(_t, "frames_to_nhwc_%dx%dx%d" % (b * t, h, w), 0.0, 18.0 * b * t * h * w, replay=call if _t is not None else None,
             keep=(video, out))
    LAUNCH_COUNTS["frames_to_nhwc_bf16_kernel"] += 1
    return out.permute(0, 3, 1, 2)  # logical (N, 3, H, W), channels-last strides


class PaddedFrames:
    """Frames in the layout the TMA stem convolution reads (fod_frames_to_nhwc4p_bf16): `buf` (N, H, W + 8, 4) bf16 with
    pixel x at padded column x + 3, zero fourth channel and zero border columns.  `nchw()` is the plain (N, 3, H, W)
    view of the same memory for every other consumer."""

    def __init__(self, buf, h, w):
        self.buf, self.h, self.w = buf, h, w
        self.shape = (buf.shape[0], 3, h, w)
        self.device, self.dtype, self.is_cuda, self.requires_grad = buf.device, buf.dtype, buf.is_cuda, False

    def nchw(self):
        return self.buf[:, :, 3:3 + self.w, :3].permute(0, 3, 1, 2)


def frames_to_nhwc4p_bf16(video, frames=None):
    """video (B, L, 3, H, W) fp32 with dense (3, H, W) frames -> PaddedFrames of the first `frames` (default L) frames of
    every clip, in one pass."""
    L = lib.load()
    b, l, c, h, w = video.shape
    t = l if frames is None else frames
    assert video.dtype == torch.float32 and c == 3 and 0 < t <= l and video.stride()[2:] == (h * w, w, 1) and w % 8 == 0
    out = torch.empty((b * t, h, w + 8, 4), dtype=BF16, device=video.device)
    _t = prof.begin()
    args = (_p(video), video.stride(0), video.stride(1), b, t, h, w, _p(out))
    call = lambda: L.fod_frames_to_nhwc4p_bf16(*args, _stream())
    lib.check(call(), "fod_frames_to_nhwc4p_bf16")
    prof.end(_t, "frames_to_nhwc4p_%dx%dx%d" % (b * t, h, w), 0.0, (12.0 * w + 8.0 * (w + 8)) * b * t * h,
             replay=call if _t is not None else None, keep=(video, out))
    LAUNCH_COUNTS["frames_to_nhwc4p_bf16_kernel"] += 1
    return PaddedFrames(out, h, w)


def relu_bwd_add(a, b, y):
    """(a [+ b]) * (y > 0) in one pass; a, b, y bf16 CUDA tensors of one shape AND one memory layout (dense)."""
    L = lib.load()
    _chk_bf16(a, "relu_bwd_add a")
    _chk_bf16(y, "relu_bwd_add y")
    assert a.shape == y.shape and a.stride() == y.stride() and (b is None or (b.shape == a.shape and b.stride() == a.stride()))
    out = torch.empty_like(a)  # preserves the (dense) strides
    assert out.stride() == a.stride() and a.numel() % 8 == 0
    _t = prof.begin()
    ptrs = (_p(a), _p(b), _p(y), _p(out), a.numel())
    call = lambda: L.fod_relu_bwd_add(*ptrs, _stream())
    lib.check(call(), "fod_relu_bwd_add")
    prof.end(_t, "relu_bwd_add%d_%d" % (3 if b is not None else 2, a.numel()), 0.0, (8.0 if b is not None else 6.0) * a.numel(),
             replay=call if _t is not None else None, keep=(a, b, y, out))
    LAUNCH_COUNTS["relu_bwd_add_kernel"] += 1
    return out


def stem_weights(w, scale):
    """(64,3,7,7) fp32 conv weight and (64) fold scale -> the (64,192) bf16 operand of fod_stem_conv7x7s2:
    [o][kh*24 + 1 + kw*3 + c] = w[o][c][kh][kw] * scale[o], zeros elsewhere."""
    assert tuple(w.shape) == (64, 3, 7, 7)
    wf = (w.detach().float() * scale.view(-1, 1, 1, 1)).permute(0, 2, 3, 1).reshape(64, 7, 21)  # [o][kh][kw*3+c]
    wk = torch.zeros((64, 7, 24), dtype=torch.float32, device=w.device)
    wk[:, :, 1:22] = wf
    return torch.cat([wk.reshape(64, 168), torch.zeros((64, 24), device=w.device)], 1).to(BF16).contiguous()


def stem_weights4(w, scale):
    """(64,3,7,7) fp32 conv weight and (64) fold scale -> the (64,224) bf16 operand of fod_stem_conv7x7s2_p4:
    [o][kh*32 + kw*4 + c] = w[o][c][kh][kw] * scale[o] (kw < 7, c < 3), zeros elsewhere."""
    assert tuple(w.shape) == (64, 3, 7, 7)
    wf = (w.detach().float() * scale.view(-1, 1, 1, 1)).permute(0, 2, 3, 1)  # [o][kh][kw][c]
    wk = torch.zeros((64, 7, 8, 4), dtype=torch.float32, device=w.device)
    wk[:, :, :7, :3] = wf
    return wk.reshape(64, 224).to(BF16).contiguous()


def stem_conv7x7s2_p4(frames, wk4, bias):
    """relu(conv7x7/s2/p3(x, W·s) + t) for PaddedFrames -> channels_last (N,64,H/2,W/2); operand tiles by TMA."""
    L = lib.load()
    buf, h, w = frames.buf, frames.h, frames.w
    _chk_bf16(buf, "stem x4p")
    _chk_bf16(wk4, "stem wk4")
    n = buf.shape[0]
    assert buf.is_contiguous() and tuple(buf.shape) == (n, h, w + 8, 4) and w % 8 == 0 and h % 2 == 0
    assert tuple(wk4.shape) == (64, 224) and wk4.is_contiguous() and bias.dtype == torch.float32 and bias.numel() == 64
    oh, ow = h // 2, w // 2
    y = torch.empty((n, 64, oh, ow), dtype=BF16, device=buf.device, memory_format=torch.channels_last)
    _t = prof.begin()
    call = lambda: L.fod_stem_conv7x7s2_p4(_p(buf), _p(wk4), _p(bias), _p(y), n, h, w, _stream())
    lib.check(call(), "fod_stem_conv7x7s2_p4")
    prof.end(_t, "stem_conv7x7s2_%dx%d" % (h, w), 2.0 * n * oh * ow * 64 * 147, 2.0 * (buf.numel() + y.numel()),
             replay=call if _t is not None else None, keep=(buf, wk4, bias, y))
    LAUNCH_COUNTS["stem_conv7x7s2_p4_kernel"] += 1
    return y


def stem_conv_pool_p4(frames, wk4, bias):
    """maxpool3x3/s2/p1(relu(conv7x7/s2/p3(x, W·s) + t)) for PaddedFrames in ONE kernel -> channels_last (N,64,PH,PW)."""
    L = lib.load()
    buf, h, w = frames.buf, frames.h, frames.w
    _chk_bf16(buf, "stem x4p")
    _chk_bf16(wk4, "stem wk4")
    n = buf.shape[0]
    assert buf.is_contiguous() and tuple(buf.shape) == (n, h, w + 8, 4) and w % 8 == 0 and h % 2 == 0
    assert tuple(wk4.shape) == (64, 224) and wk4.is_contiguous() and bias.dtype == torch.float32 and bias.numel() == 64
    oh, ow = h // 2, w // 2
    ph, pw = (oh - 1) // 2 + 1, (ow - 1) // 2 + 1
    y = torch.empty((n, 64, ph, pw), dtype=BF16, device=buf.device, memory_format=torch.channels_last)
    _t = prof.begin()
    call = lambda: L.fod_stem_conv_pool_p4(_p(buf), _p(wk4), _p(bias), _p(y), n, h, w, _stream())
    lib.check(call(), "fod_stem_conv_pool_p4")
    prof.end(_t, "stem_conv_pool_%dx%d" % (h, w), 2.0 * n * oh * ow * 64 * 147, 2.0 * (buf.numel() + y.numel()),
             replay=call if _t is not None else None, keep=(buf, wk4, bias, y))
    LAUNCH_COUNTS["stem_conv_pool_p4_kernel"] += 1
    return y


def stem_conv7x7s2(x, wk, bias):
    """relu(conv7x7/s2/p3(x, W·s) + t) for a channels_last bf16 (N,3,H,W) image batch -> channels_last (N,64,OH,OW)."""
    L = lib.load()
    _chk_bf16(x, "stem x")
    _chk_bf16(wk, "stem wk")
    n, c, h, w = x.shape
    assert c == 3 and x.is_contiguous(memory_format=torch.channels_last) and w % 8 == 0
    assert tuple(wk.shape) == (64, 192) and wk.is_contiguous() and bias.dtype == torch.float32 and bias.numel() == 64
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.empty((n, 64, oh, ow), dtype=BF16, device=x.device, memory_format=torch.channels_last)
    _t = prof.begin()
    lib.check(L.fod_stem_conv7x7s2(_p(x), _p(wk), _p(bias), _p(y), n, h, w, _stream()), "fod_stem_conv7x7s2")
    prof.end(_t, "stem_conv7x7s2_%dx%d" % (h, w), 2.0 * n * oh * ow * 64 * 147, 2.0 * (x.numel() + y.numel()),
             replay=(lambda: L.fod_stem_conv7x7s2(_p(x), _p(wk), _p(bias), _p(y), n, h, w, _stream())) if _t is not None else None,
             keep=(x, wk, bias, y))
    LAUNCH_COUNTS["stem_conv7x7s2_kernel"] += 1
    return y


class ConvFoldTable:
    """Descriptor table (device memory) + persistent bf16 channels_last outputs for fod_fold_conv_weights: every
    trainable backbone conv's W*s in ONE launch per step.  weights: fp32 (O,I,KH,KW) parameters; scales: fp32 (O)."""

    def __init__(self, weights, scales):
        dev = weights[0].device
        self.weights, self.scales = list(weights), list(scales)
        self.outs = [torch.empty(w.shape, dtype=BF16, device=dev).contiguous(memory_format=torch.channels_last) for w in weights]
        arr = (lib.FoldDesc * len(weights))()
        self.max_elems = 0
        for d, w, s, o in zip(arr, self.weights, self.scales, self.outs):
            assert w.dtype == torch.float32 and w.is_contiguous() and s.dtype == torch.float32 and s.numel() == w.shape[0]
            O, I, KH, KW = w.shape
            assert KH * KW > 1 or I % 4 == 0
            assert o.permute(0, 2, 3, 1).is_contiguous()
            d.w, d.scale, d.out, d.O, d.I, d.HW = w.data_ptr(), s.data_ptr(), o.data_ptr(), O, I, KH * KW
            self.max_elems = max(self.max_elems, w.numel())
        raw = bytes(arr)
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.key = tuple((w.data_ptr(), s.data_ptr()) for w, s in zip(self.weights, self.scales))
        self.bytes = sum(6.0 * w.numel() for w in weights)

    def valid_for(self, weights, scales):
        return self.key == tuple((w.data_ptr(), s.data_ptr()) for w, s in zip(weights, scales))

    def run(self):
        L = lib.load()
        _t = prof.begin()
        ptrs = (_p(self.table), len(self.outs), self.max_elems)
        call = lambda: L.fod_fold_conv_weights(*ptrs, _stream())
        lib.check(call(), "fod_fold_conv_weights")
        prof.end(_t, "fold_conv_weights", 0.0, self.bytes, replay=call if _t is not None else None, keep=(self.table,))
        LAUNCH_COUNTS["fold_conv_weights_kernel"] += 1
        return self.outs


def conv_wgrad_unfold(dwq, scale, grad):
    """grad (O,I,KH,KW) fp32 contiguous += float(dwq) * scale[o]; dwq bf16 (O,I,KH,KW) in channels_last storage."""
    L = lib.load()
    O, I, KH, KW = grad.shape
    assert dwq.dtype == BF16 and dwq.shape == grad.shape and dwq.permute(0, 2, 3, 1).is_contiguous()
    assert grad.dtype == torch.float32 and grad.is_contiguous() and scale.dtype == torch.float32
    _t = prof.begin()
    ptrs = (_p(dwq), _p(scale), _p(grad), O, I, KH * KW)
    call = lambda: L.fod_conv_wgrad_unfold(*ptrs, _stream())
    lib.check(call(), "fod_conv_wgrad_unfold")
    prof.end(_t, "conv_wgrad_unfold_%dx%dx%d" % (O, I, KH * KW), 0.0, 10.0 * grad.numel(),
             replay=call if _t is not None else None, keep=(dwq, scale, grad))
    LAUNCH_COUNTS["conv_wgrad_unfold_kernel"] += 1


def targets_compact(boxes_xyxy, classes, active, height, width):
    """st_detr.py:237-263 on the device: returns (tgt_boxes (B,slots,4) cxcywh, tgt_labels int32, tgt_count int32,
    total_count fp32 scalar tensor)."""
    L = lib.load()
    B, slots = active.shape
    dev = boxes_xyxy.device
    boxes_xyxy = boxes_xyxy.contiguous().float()
    classes = classes.contiguous().long()
    active = active.contiguous().long()
    ob = torch.zeros((B, slots, 4), dtype=torch.float32, device=dev)
    ol = torch.zeros((B, slots), dtype=torch.int32, device=dev)
    oc = torch.empty((B,), dtype=torch.int32, device=dev)
    tot = torch.zeros((1,), dtype=torch.float32, device=dev)
    inv_w = float(torch.tensor(1 / width, dtype=torch.float32))
    inv_h = float(torch.tensor(1 / height, dtype=torch.float32))
    lib.check(L.fod_targets_compact(_p(boxes_xyxy), _p(classes), _p(active), _p(ob), _p(ol), _p(oc), _p(tot), B, slots,
                                    inv_w, inv_h, _stream()), "fod_targets_compact")
    LAUNCH_COUNTS["targets_compact_kernel"] += 1
    return ob, ol, oc, tot


def match(logits, boxes, tgt_boxes, tgt_labels, tgt_count, w_class=2.0, w_bbox=5.0, w_giou=2.0, alpha=0.25,
          want_lists=True, want_cost=False, status=None):
    """All-level Hungarian matching (matcher.py:40-87).  logits (L,B,M,C) fp32, boxes (L,B,M,4) fp32.

    Returns dict(match_tgt (L,B,M) int32, pred_idx/tgt_idx (L,B,KMAX) int64 padded with -1, status (1,) int32, cost)."""
    Lb = lib.load()
    logits = logits.contiguous().float()
    boxes = boxes.contiguous().float()
    Lv, B, M, Cn = logits.shape
    NT = tgt_boxes.shape[1]
    dev = logits.device
    kmax = min(M, NT)
    out = {"match_tgt": torch.empty((Lv, B, M), dtype=torch.int32, device=dev),
           "status": status if status is not None else torch.zeros((1,), dtype=torch.int32, device=dev)}
    a = lib.MatchArgs()
    a.logits, a.boxes, a.tgt_boxes, a.tgt_labels, a.tgt_count = _p(logits), _p(boxes), _p(tgt_boxes), _p(tgt_labels), _p(tgt_count)
    a.match_tgt, a.status = _p(out["match_tgt"]), _p(out["status"])
    if want_lists:
        out["pred_idx"] = torch.empty((Lv, B, kmax), dtype=torch.int64, device=dev)
        out["tgt_idx"] = torch.empty((Lv, B, kmax), dtype=torch.int64, device=dev)
        a.pred_idx, a.tgt_idx = _p(out["pred_idx"]), _p(out["tgt_idx"])
    if want_cost:
        out["cost"] = torch.zeros((Lv, B, M, NT), dtype=torch.float32, device=dev)
        a.cost_out = _p(out["cost"])
    if M * NT * 4 + 24 * (M + NT) + M * Cn * 4 + 64 > 200 * 1024:
        out["_ws"] = torch.empty((Lv, B, M, NT), dtype=torch.float32, device=dev)
        a.cost_ws = _p(out["_ws"])
    a.L, a.B, a.M, a.C, a.NT = Lv, B, M, Cn, NT
    a.w_class, a.w_bbox, a.w_giou, a.alpha = w_class, w_bbox, w_giou, alpha
    _t = prof.begin()
    lib.check(Lb.fod_match(C.byref(a), _stream()), "fod_match")
    prof.end(_t, "match", 0.0, 4.0 * Lv * B * M * (Cn + 4 + 1),
             replay=(lambda: Lb.fod_match(C.byref(a), _stream())) if _t is not None else None,
             keep=(logits, boxes, tgt_boxes, tgt_labels, tgt_count, out))
    LAUNCH_COUNTS["match_kernel"] += 1
    return out


def set_loss(logits, boxes, match_tgt, tgt_boxes, tgt_labels, tgt_count, num_boxes, w_ce=2.0, w_bbox=5.0, w_giou=2.0,
             alpha=0.25, want_grads=True):
    """Losses + their gradients (set_criterion.py:36-115).  Returns (out (8+3L,) fp32, dlogits, dboxes)."""
    Lb = lib.load()
    logits = logits.contiguous().float()
    boxes = boxes.contiguous().float()
    Lv, B, M, Cn = logits.shape
    dev = logits.device
    out = torch.zeros((8 + 3 * Lv,), dtype=torch.float32, device=dev)
    partial = torch.empty((Lv, B, 8), dtype=torch.float32, device=dev)
    dl = torch.empty_like(logits) if want_grads else None
    db = torch.empty_like(boxes) if want_grads else None
    a = lib.SetLossArgs()
    a.logits, a.boxes, a.match_tgt = _p(logits), _p(boxes), _p(match_tgt)
    a.tgt_boxes, a.tgt_labels, a.tgt_count, a.num_boxes = _p(tgt_boxes), _p(tgt_labels), _p(tgt_count), _p(num_boxes)
    a.dlogits, a.dboxes, a.partial, a.out = _p(dl), _p(db), _p(partial), _p(out)
    a.L, a.B, a.M, a.C, a.NT, a.main_level = Lv, B, M, Cn, tgt_boxes.shape[1], Lv - 1
    a.w_ce, a.w_bbox, a.w_giou, a.alpha = w_ce, w_bbox, w_giou, alpha
    _t = prof.begin()
    lib.check(Lb.fod_set_loss(C.byref(a), _stream()), "fod_set_loss")
    prof.end(_t, "set_loss", 0.0, 8.0 * Lv * B * M * (Cn + 4),
             replay=(lambda: Lb.fod_set_loss(C.byref(a), _stream())) if _t is not None else None,
             keep=(logits, boxes, match_tgt, tgt_boxes, tgt_labels, tgt_count, num_boxes, dl, db, partial, out))
    LAUNCH_COUNTS["set_loss_kernel"] += 1
    LAUNCH_COUNTS["set_loss_finalize_kernel"] += 1
    return out, dl, db


_AP_THR = {}


def ap_match(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk=50):
    """utils/od_map.py:214-287 in one launch.  pred_boxes (B,M,4) xyxy px fp32, scores (B,M,C) with the generic class
    last, annotation slots (B,N,...).  Returns (confs (10,C,B*K) fp32, is_positive bool, size_categories (C,4,B*K)
    bool, num_annos (C,4) int64)."""
    L = lib.load()
    dev = pred_boxes.device
    B, M, Cn = scores.shape
    N = anno_boxes.shape[1]
    K = min(topk, M)
    key = str(dev)
    if key not in _AP_THR:  # od_map.py:24: the ten IoU thresholds, torch.arange's own fp32 values
        _AP_THR[key] = torch.arange(0.50, 1.00, 0.05).to(dev)
    thr = _AP_THR[key]
    T = thr.numel()
    pb = pred_boxes.contiguous().float()
    sc = scores.contiguous().float()
    ab = anno_boxes.contiguous().float()
    ac = anno_classes.contiguous().long()
    aa = anno_active.contiguous().long()
    confs = torch.empty((T, Cn, B * K), dtype=torch.float32, device=dev)
    is_pos = torch.empty((T, Cn, B * K), dtype=torch.bool, device=dev)
    size_cat = torch.empty((Cn, 4, B * K), dtype=torch.bool, device=dev)
    num_annos = torch.zeros((Cn, 4), dtype=torch.int64, device=dev)
    h, w = imsize
    a = lib.ApArgs()
    a.pred_boxes, a.scores, a.anno_boxes, a.anno_classes, a.anno_active = _p(pb), _p(sc), _p(ab), _p(ac), _p(aa)
    a.thresholds, a.confs, a.is_positive, a.size_cat, a.num_annos = _p(thr), _p(confs), _p(is_pos), _p(size_cat), _p(num_annos)
    a.B, a.M, a.C, a.N, a.K, a.T = B, M, Cn, N, K, T
    # od_map.py:20-23 delimiters times H*W, evaluated as the reference does (python floats -> fp32 at the comparison)
    a.size_lo, a.size_hi = (1 / 24) * (1 / 64) * h * w, (1 / 4) * (1 / 12) * h * w
    _t = prof.begin()
    lib.check(L.fod_ap_match(C.byref(a), _stream()), "fod_ap_match")
    prof.end(_t, "ap_match", 0.0, 4.0 * B * (M * (Cn + 4) + N * 6),
             replay=(lambda: L.fod_ap_match(C.byref(a), _stream())) if _t is not None else None,
             keep=(pb, sc, ab, ac, aa, thr, confs, is_pos, size_cat, num_annos))
    LAUNCH_COUNTS["ap_match_kernel"] += 1
    return confs, is_pos, size_cat, num_annos
