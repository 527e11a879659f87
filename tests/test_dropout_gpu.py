"""Train-mode dropout (reference: p = 0.1 on the attention weights and after every sub-block, transformer.py:62-64,
92-103,201-234,271-310,403-417; torch.nn.MultiheadAttention / ConditionalDETR attention.py:370).

The masks come from a counter-based generator inside the kernels (csrc/dropout.cuh), so the reference's RNG stream
cannot be matched; parity is
  * exact at p = 0 (every other parity test),
  * exact against a torch fp32 reference that applies THE SAME mask — the mask is a function of (seed, step, call id,
    element index) only, so it is read back from a probe launch (identity operand) and handed to torch,
  * statistical: keep rate, 1/(1-p) scaling, fresh masks per step / call site, identical masks forward and backward.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
P = 0.1


def _dev():
    return torch.device("cuda:0")


def _rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(BF).to(_dev())


def _rel(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / max(1e-30, float(b.norm())))


def _gemm_mask(M, N, drop):
    """Keep mask (M, N) of a GEMM-epilogue dropout site: read back from a probe launch whose product is 1 everywhere."""
    from fod_b200 import ops
    x = torch.ones(M, 256, dtype=BF, device=_dev())
    w = torch.full((N, 256), 1.0 / 256, dtype=BF, device=_dev())
    y = ops.gemm(x, w, drop=drop).float()
    keep = y > 0.5
    assert torch.allclose(y[keep], torch.full_like(y[keep], 1.0 / (1.0 - drop[0])), rtol=1e-2)
    assert float(y[~keep].abs().max() if (~keep).any() else 0.0) == 0.0
    return keep


@pytest.mark.parametrize("M,N", [(5600, 2048), (300, 256), (77, 768), (22400, 2048)])
def test_gemm_dropout_rate_scaling_and_streams(M, N):
    from fod_b200 import ops
    ops.set_dropout_seed(1234, _dev())
    keep = _gemm_mask(M, N, (P, 7))
    rate = float(keep.float().mean())
    sigma = (P * (1 - P) / keep.numel()) ** 0.5
    assert abs(rate - (1 - P)) < 5 * sigma + 2e-5, (rate, sigma)   # threshold is round(p * 65536): bias < 1e-5
    # rows and columns are unbiased too (no striping from the 8-lane grouping)
    assert float((keep.float().mean(0) - (1 - P)).abs().max()) < 6 * (P * (1 - P) / M) ** 0.5 + 1e-3
    assert float((keep.float().mean(1) - (1 - P)).abs().max()) < 6 * (P * (1 - P) / N) ** 0.5 + 1e-3
    # same (seed, step, call) -> same mask; another call id, step or seed -> an independent one
    assert torch.equal(keep, _gemm_mask(M, N, (P, 7)))
    other = _gemm_mask(M, N, (P, 8))
    agree = float((keep == other).float().mean())
    assert abs(agree - (P * P + (1 - P) ** 2)) < 0.01, agree
    ops.rng_advance(_dev())
    agree = float((keep == _gemm_mask(M, N, (P, 7))).float().mean())
    assert abs(agree - (P * P + (1 - P) ** 2)) < 0.01, agree
    ops.set_dropout_seed(1235, _dev())
    agree = float((keep == _gemm_mask(M, N, (P, 7))).float().mean())
    assert abs(agree - (P * P + (1 - P) ** 2)) < 0.01, agree
    # applied after bias and ReLU
    x, w, b = _rnd(M, 256), _rnd(N, 256, scale=0.06), torch.randn(N, device=_dev())
    k2 = _gemm_mask(M, N, (P, 3))
    ref = (x.float() @ w.float().t() + b).relu() * k2 / (1 - P)
    assert _rel(ops.gemm(x, w, bias=b, relu=True, drop=(P, 3)), ref) < 2e-2


@pytest.mark.parametrize("M", [5600, 300, 16])
def test_linear_ln_dropout_matches_torch_with_same_mask(M):
    """y = LN(res + dropout(x W^T + b)) forward and every gradient (transformer.py:271,285,417)."""
    from fod_b200 import autograd as A
    from fod_b200 import ops
    ops.set_dropout_seed(99, _dev())
    D = 256
    drop = (P, 5)
    keep = _gemm_mask(M, D, drop).float() / (1 - P)
    x, res = _rnd(M, D, seed=1), _rnd(M, D, seed=2)
    w = (torch.randn(D, D, device=_dev()) * 0.06).requires_grad_()
    b = torch.randn(D, device=_dev()).requires_grad_()
    gam = (torch.rand(D, device=_dev()) + 0.5).requires_grad_()
    bet = torch.randn(D, device=_dev()).requires_grad_()
    xg, rg = x.clone().requires_grad_(), res.clone().requires_grad_()
    y = A.linear_ln(xg, rg, w, b, w.detach().to(BF), gam, bet, 1e-5, drop=drop)
    dy = _rnd(M, D, seed=3)
    y.backward(dy)
    got = [y, xg.grad, rg.grad, w.grad, b.grad, gam.grad, bet.grad]
    x32, r32 = x.float().requires_grad_(), res.float().requires_grad_()
    w32, b32 = w.detach().to(BF).float().requires_grad_(), b.detach().clone().requires_grad_()
    g32, be32 = gam.detach().clone().requires_grad_(), bet.detach().clone().requires_grad_()
    u = (x32 @ w32.t() + b32) * keep
    yr = torch.nn.functional.layer_norm(r32 + u, (D,), g32, be32, 1e-5)
    yr.backward(dy.float())
    ref = [yr, x32.grad, r32.grad, w32.grad, b32.grad, g32.grad, be32.grad]
    for name, a, r in zip(("y", "dx", "dres", "dw", "db", "dgamma", "dbeta"), got, ref):
        assert _rel(a, r) < 2e-2, (name, _rel(a, r))


@pytest.mark.parametrize("M", [5600, 128])
def test_ffn_ln_dropout_matches_torch_with_same_masks(M):
    """y = LN(x + dropout(W2 dropout(relu(W1 x + b1)) + b2)) (transformer.py:407-419, 228-235, 310)."""
    from fod_b200 import autograd as A
    from fod_b200 import ops
    ops.set_dropout_seed(5, _dev())
    D, F = 256, 2048
    dh, do = (P, 11), (P, 12)
    kh = _gemm_mask(M, F, dh).float() / (1 - P)
    ko = _gemm_mask(M, D, do).float() / (1 - P)
    x = _rnd(M, D, seed=1)
    w1 = (torch.randn(F, D, device=_dev()) * 0.06).requires_grad_()
    b1 = (torch.randn(F, device=_dev()) * 0.1).requires_grad_()
    w2 = (torch.randn(D, F, device=_dev()) * 0.03).requires_grad_()
    b2 = (torch.randn(D, device=_dev()) * 0.1).requires_grad_()
    gam = (torch.rand(D, device=_dev()) + 0.5).requires_grad_()
    bet = torch.randn(D, device=_dev()).requires_grad_()
    xg = x.clone().requires_grad_()
    y = A.ffn_ln(xg, w1, b1, w1.detach().to(BF), w2, b2, w2.detach().to(BF), gam, bet, 1e-5, drop_h=dh, drop_o=do)
    dy = _rnd(M, D, seed=3)
    y.backward(dy)
    got = [y, xg.grad, w1.grad, b1.grad, w2.grad, b2.grad, gam.grad, bet.grad]
    x32 = x.float().requires_grad_()
    p = [t.detach().to(BF).float().requires_grad_() if t.dim() == 2 else t.detach().clone().requires_grad_()
         for t in (w1, b1, w2, b2, gam, bet)]
    hr = (x32 @ p[0].t() + p[1]).relu() * kh
    yr = torch.nn.functional.layer_norm(x32 + (hr @ p[2].t() + p[3]) * ko, (D,), p[4], p[5], 1e-5)
    yr.backward(dy.float())
    ref = [yr, x32.grad] + [t.grad for t in p]
    for name, a, r in zip(("y", "dx", "dw1", "db1", "dw2", "db2", "dgamma", "dbeta"), got, ref):
        assert _rel(a, r) < 2e-2, (name, _rel(a, r))


def test_backward_uses_the_masks_of_its_own_forward_pass():
    """Two forward passes before the first backward (gradient accumulation): every dropout site carries a snapshot of
    {seed, step} of ITS pass, so the first pass's backward regenerates the first pass's masks."""
    from fod_b200 import autograd as A
    from fod_b200 import ops
    D, M = 256, 600
    x, res = _rnd(M, D, seed=1), _rnd(M, D, seed=2)
    w = (torch.randn(D, D, device=_dev()) * 0.06)
    b, gam, bet = torch.randn(D, device=_dev()), torch.rand(D, device=_dev()) + 0.5, torch.randn(D, device=_dev())
    dy = _rnd(M, D, seed=3)

    def forward():
        ops.begin_forward()
        xg = x.clone().requires_grad_()
        y = A.linear_ln(xg, res, w, b, w.to(BF), gam, bet, 1e-5, drop=ops.drop_site(P, True))
        return xg, y

    ops.set_dropout_seed(321, _dev())
    xg, y = forward()
    y.backward(dy)
    y_ref, g_ref = y.detach().clone(), xg.grad.clone()
    ops.set_dropout_seed(321, _dev())
    xg1, y1 = forward()
    xg2, y2 = forward()                      # advances the step: other masks
    assert torch.equal(y1, y_ref) and not torch.equal(y2, y_ref)
    y1.backward(dy)                           # after the second forward
    assert torch.equal(xg1.grad, g_ref)
    y2.backward(dy)
    assert not torch.equal(xg2.grad, g_ref)


def _attn_mask(B, H, Lq, Lk, scale, q1, k1, drop, q2=None, k2=None):
    """Dropped-out attention weights read back through one-hot values, 32 keys per probe launch: (B, H, Lq, Lk)."""
    from fod_b200 import ops
    out = torch.zeros(B, H, Lq, Lk, device=_dev())
    for c0 in range(0, Lk, 32):
        v = torch.zeros(B, Lk, H, 32, dtype=BF, device=_dev())
        n = min(32, Lk - c0)
        idx = torch.arange(n, device=_dev())
        v[:, c0 + idx, :, idx] = 1.0
        o, _ = ops.attn_fwd(q1, k1, v.view(B, Lk, H * 32), scale, H, q2=q2, k2=k2, drop=drop)
        out[..., c0:c0 + n] = o.view(B, Lq, H, 32).permute(0, 2, 1, 3)[..., :n].float()
    return out


def _softmax_ref(q, k, scale, H):
    B, Lq, C = q.shape
    dh = C // H
    qh = q.float().view(B, Lq, H, dh).transpose(1, 2)
    kh = k.float().view(B, -1, H, dh).transpose(1, 2)
    return torch.softmax(qh @ kh.transpose(2, 3) * scale, dim=-1)


@pytest.mark.parametrize("B,Lq,Lk", [(3, 350, 350), (2, 128, 350), (2, 300, 7), (4, 350, 1), (1, 700, 700), (1, 1400, 1400)])
def test_attention_dropout_forward_and_backward_share_the_mask(B, Lq, Lk):
    """Softmax weights are dropped AFTER normalisation (F.dropout(softmax(..)), attention.py:370); forward and both
    backward paths (two-kernel and fused) regenerate the identical mask."""
    from fod_b200 import ops
    ops.set_dropout_seed(77, _dev())
    H, scale, drop = 8, 32 ** -0.5, (P, 21)
    q, k, v = _rnd(B, Lq, 256, seed=1), _rnd(B, Lk, 256, seed=2), _rnd(B, Lk, 256, seed=3)
    pd = _attn_mask(B, H, Lq, Lk, scale, q, k, drop)               # dropout(softmax) as the kernel applies it
    p = _softmax_ref(q, k, scale, H)
    keep = pd > 0.5 * p
    rate = float(keep.float().mean())
    sigma = (P * (1 - P) / keep.numel()) ** 0.5
    assert abs(rate - (1 - P)) < 5 * sigma + 2e-5, (rate, sigma)
    big = p > 1e-3
    assert _rel(pd[big & keep], p[big & keep] / (1 - P)) < 2e-2
    assert float(pd[~keep].abs().max() if (~keep).any() else 0.0) == 0.0
    # forward output and gradients against torch with this mask
    mask = keep.float() / (1 - P)
    out, lse = ops.attn_fwd(q, k, v, scale, H, drop=drop)
    q32, k32, v32 = q.float().requires_grad_(), k.float().requires_grad_(), v.float().requires_grad_()
    pr = _softmax_ref(q32, k32, scale, H) * mask
    ref = (pr @ v32.view(B, Lk, H, 32).transpose(1, 2)).transpose(1, 2).reshape(B, Lq, 256)
    assert _rel(out, ref) < 2e-2
    dout = _rnd(B, Lq, 256, seed=4)
    ref.backward(dout.float())
    modes = [1] + ([2] if ops.attn_bwd_is_fused(Lq, False) else [])
    for mode in modes:
        dq, dk, dv, _, _ = ops.attn_bwd(q, k, v, out, lse, dout, scale, H, mode=mode, drop=drop)
        for name, a, r in (("dq", dq, q32.grad), ("dk", dk, k32.grad), ("dv", dv, v32.grad)):
            if Lk == 1 and name != "dv":
                # one key: the softmax weight is 1 whatever q and k are, their true gradients are 0; what the kernels
                # produce is the bf16 rounding of O inside dS = P (dP - rowsum(dO O)) — small against dV
                assert float(a.float().norm()) < 2e-2 * float(dv.float().norm()), (mode, name)
                continue
            assert _rel(a, r) < 2.5e-2, (mode, name, _rel(a, r))


@pytest.mark.parametrize("B,M,S", [(2, 128, 350), (1, 128, 1400)])
def test_cross_attention_dropout_two_part_heads(B, M, S):
    """Decoder cross-attention (head dim 64 = content ‖ positional part): same mask in forward and backward."""
    from fod_b200 import autograd as A
    from fod_b200 import ops
    ops.set_dropout_seed(31, _dev())
    H, scale, drop = 8, 64 ** -0.5, (P, 4)
    qc, qs = _rnd(B, M, 256, seed=1), _rnd(M, 256, seed=2)
    kv, kp = _rnd(B, S, 512, seed=3), _rnd(S, 256, seed=4)
    pd = _attn_mask(B, H, M, S, scale, qc, kv[..., :256], drop, q2=qs, k2=kp)

    def heads(t, n):
        return t.float().view(-1, n, H, 32).transpose(1, 2)

    def weights(qc_, qs_, k_, kp_):
        s = heads(qc_, M) @ heads(k_, S).transpose(2, 3) + heads(qs_, M) @ heads(kp_, S).transpose(2, 3)
        return torch.softmax(s * scale, dim=-1)

    p = weights(qc, qs, kv[..., :256], kp)
    keep = pd > 0.5 * p
    assert abs(float(keep.float().mean()) - (1 - P)) < 5 * (P * (1 - P) / keep.numel()) ** 0.5 + 2e-5
    mask = keep.float() / (1 - P)
    qcg, qsg, kvg, kpg = (t.clone().requires_grad_() for t in (qc, qs, kv, kp))
    out = A.cross_attention(qcg, qsg, kvg, kpg, H, scale, drop=drop)
    dout = _rnd(B, M, 256, seed=5)
    out.backward(dout)
    r = [t.float().requires_grad_() for t in (qc, qs, kv, kp)]
    pr = weights(r[0], r[1], r[2][..., :256], r[3]) * mask
    ref = (pr @ heads(r[2][..., 256:], S)).transpose(1, 2).reshape(B, M, 256)
    ref.backward(dout.float())
    assert _rel(out, ref) < 2e-2
    for name, a, b in zip(("dqc", "dqs", "dkv", "dkp"), (qcg.grad, qsg.grad, kvg.grad, kpg.grad), [t.grad for t in r]):
        assert _rel(a, b) < 2.5e-2, (name, _rel(a, b))


def test_model_dropout_statistics_and_graph_replay(ref_state_dict):
    """Whole model: p = 0 in train mode equals eval exactly; p = 0.1 gives finite losses/gradients that differ from
    step to step (also across replays of ONE captured graph), and whose mean stays near the dropout-free loss."""
    from argparse import Namespace

    from fod_b200 import SpatioTemporalDETRArgs, build_model, ops, set_dropout
    from fod_b200 import train as T
    from fod_b200.synthetic import make_clip
    dev = _dev()
    dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
    model = build_model(Namespace(device=dev, distributed=False), dargs)          # reference default: p = 0.1
    model.load_state_dict(ref_state_dict)
    data = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in make_clip(2, 3, 128, 160, seed=3, fixed_targets=5).items()}
    ops.set_dropout_seed(2024, dev)
    model.eval()
    with torch.no_grad():
        l_eval = float(model(data)[2])
    set_dropout(model, 0.0)
    model.train()
    with torch.no_grad():
        assert abs(float(model(data)[2]) - l_eval) < 1e-3 * abs(l_eval)
    assert set_dropout(model, P) > 0
    losses = []
    for _ in range(12):
        model.zero_grad(set_to_none=True)
        loss = model(data)[2]
        loss.backward()
        losses.append(float(loss))
    assert all(v == v and abs(v) < 1e6 for v in losses), losses
    assert len(set(losses)) == len(losses), losses                     # fresh masks every step
    missing = [k for k, p in model.named_parameters() if p.requires_grad and p.grad is None]
    assert not missing, missing[:5]
    assert all(bool(torch.isfinite(p.grad).all()) for p in model.parameters() if p.grad is not None)
    mean = sum(losses) / len(losses)
    assert abs(mean - l_eval) < 0.35 * abs(l_eval), (mean, l_eval)
    # egodeep query/key parameters still cannot influence the output (one IMU key per sequence: softmax weight 1)
    sd = dict(model.named_parameters())
    dead = [k for k in sd if ".egodeep_attend.query_content." in k or ".egodeep_attend.key." in k]
    assert dead and all(float(sd[k].grad.abs().max()) == 0.0 for k in dead)
    # one captured graph, replayed: the device-side step counter advances inside the graph
    del loss  # a live eager autograd graph pins the parameters' AccumulateGrad nodes to the eager (legacy) stream
    model.zero_grad(set_to_none=True)
    opt = T.setup_optimizer(model, dargs, capturable=True)
    for g in opt.param_groups:
        g["lr"] = 0.0
        g["weight_decay"] = 0.0
    step = T.GraphedTrainStep(model, opt, data, dargs.max_norm, warmup=1)
    replay = [float(step(data)) for _ in range(6)]
    assert len(set(replay)) == len(replay), replay
    assert all(abs(v - l_eval) < 0.6 * abs(l_eval) for v in replay), (replay, l_eval)


def test_joint_mode_trains_with_dropout(ref_state_dict):
    """JointEncoder mode with dropout on: the egodeep block attends over T IMU keys per token (PlainAttnFn with dropout
    on the weights, Lk = T) and the space-time self-attention runs at S = T*h*w: finite loss, every trainable parameter
    gets a finite gradient, and two steps differ."""
    from argparse import Namespace

    from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
    from fod_b200.st_detr import JointEncoder
    from fod_b200.synthetic import make_clip
    dev = _dev()
    dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
    model = build_model(Namespace(device=dev, distributed=False), dargs)   # p = 0.1
    model.load_state_dict(ref_state_dict)
    core = model._model
    core.joint_encoder = JointEncoder(core.separate_encoder.transformer)
    data = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in make_clip(2, 4, 96, 128, seed=5, fixed_targets=4).items()}
    ops.set_dropout_seed(7, dev)
    model.train()
    losses = []
    for _ in range(2):
        model.zero_grad(set_to_none=True)
        loss = model(data)[2]
        loss.backward()
        losses.append(float(loss.detach()))
    assert all(v == v and abs(v) < 1e6 for v in losses) and losses[0] != losses[1], losses
    bad = [k for k, p in model.named_parameters() if p.requires_grad and (p.grad is None or not bool(torch.isfinite(p.grad).all()))]
    assert not bad, bad[:5]
    # with several keys the egodeep query / key projections are live
    g = dict(model.named_parameters())["_model.separate_encoder.transformer.layers.0.egodeep_attend.key.weight"].grad
    assert float(g.abs().max()) > 0


@pytest.mark.parametrize("M,same", [(5600, True), (700, False)])
def test_add_ln_dropout_matches_torch_with_same_mask(M, same):
    """y = LN(x + dropout(x2)) (transformer.py:117 with x2 = x, :485-486) forward and gradients of x, x2, gamma, beta."""
    from fod_b200 import autograd as A
    from fod_b200 import ops
    ops.set_dropout_seed(17, _dev())
    D, drop = 256, (P, 9)
    keep = _gemm_mask(M, D, drop).float() / (1 - P)          # same element indexing: row * 256 + col
    x = _rnd(M, D, seed=1)
    x2 = x if same else _rnd(M, D, seed=2)
    gam = (torch.rand(D, device=_dev()) + 0.5).requires_grad_()
    bet = torch.randn(D, device=_dev()).requires_grad_()
    xg = x.clone().requires_grad_()
    x2g = xg if same else x2.clone().requires_grad_()
    y = A.add_ln_drop(xg, x2g, gam, bet, 1e-5, drop=drop)
    dy = _rnd(M, D, seed=3)
    y.backward(dy)
    a32 = x.float().requires_grad_()
    b32 = a32 if same else x2.float().requires_grad_()
    g32, be32 = gam.detach().clone().requires_grad_(), bet.detach().clone().requires_grad_()
    yr = torch.nn.functional.layer_norm(a32 + b32 * keep, (D,), g32, be32, 1e-5)
    yr.backward(dy.float())
    pairs = [("y", y, yr), ("dx", xg.grad, a32.grad), ("dgamma", gam.grad, g32.grad), ("dbeta", bet.grad, be32.grad)]
    if not same:
        pairs.append(("dx2", x2g.grad, b32.grad))
    for name, a, r in pairs:
        assert _rel(a, r) < 2e-2, (name, _rel(a, r))


def test_egodeep_head_dropout_table_equals_the_per_token_projection():
    """One IMU key per sequence, dropout on the (single) attention weight of every (token, head): the per-token
    out_proj is computed from an (n_seq, H, D) table of per-head contributions; it must equal out_proj applied to the
    value row with the dropped heads zeroed, forward and in the gradients of value rows and out_proj parameters."""
    from fod_b200.model import EgodeepAttention
    torch.manual_seed(3)
    n_seq, seq_len, D, H = 6, 50, 256, 8
    eg = EgodeepAttention(D, H, Dff=None, dropout=P).to(_dev())
    v = torch.randn(n_seq, D, device=_dev()).to(BF).requires_grad_()
    keep = (torch.rand(n_seq, seq_len, H, device=_dev()) >= P).to(BF) / (1 - P)
    o = eg._out_proj_of_masked_heads(v, keep)
    g = _rnd(n_seq * seq_len, D, seed=4)
    o.backward(g)
    got = [o, v.grad, eg.fun.out_proj.weight.grad, eg.fun.out_proj.bias.grad]
    v32 = v.detach().float().requires_grad_()
    W = eg.fun.out_proj.weight.detach().clone().requires_grad_()
    b = eg.fun.out_proj.bias.detach().clone().requires_grad_()
    a = (v32.view(n_seq, 1, H, D // H) * keep.float().unsqueeze(-1)).reshape(n_seq * seq_len, D)
    ref = a @ W.t() + b
    ref.backward(g.float())
    for name, x, r in zip(("o", "dv", "dW", "db"), got, (ref, v32.grad, W.grad, b.grad)):
        assert _rel(x, r) < 2e-2, (name, _rel(x, r))


@pytest.mark.parametrize("B,Lq,Lk,two", [(3, 350, 350, False), (2, 1400, 1400, False), (4, 128, 350, True), (2, 37, 129, False)])
def test_attention_dropout_mask_bits_are_a_cache_of_the_drawn_masks(B, Lq, Lk, two):
    """With a {seed, step} snapshot in the drop tuple the keep decisions are drawn once (fod_attn_keep_bits) and read by
    the forward and the fused backward kernel; without one every kernel draws them itself.  Same function of (seed,
    step, call, element) either way: outputs and gradients must be bit-identical, ragged key tails included."""
    from fod_b200 import ops
    ops.set_dropout_seed(4242, _dev())
    H, D = 8, 256
    q1, k1, v = _rnd(B, Lq, D, seed=1), _rnd(B, Lk, D, seed=2), _rnd(B, Lk, D, seed=3)
    q2, k2 = (_rnd(B, Lq, D, seed=4), _rnd(Lk, D, seed=5)) if two else (None, None)
    scale = (64 if two else 32) ** -0.5
    snap = ops.rng_state(_dev()).clone()
    drawn, cached = (P, 11), (P, 11, snap)
    o1, l1 = ops.attn_fwd(q1, k1, v, scale, H, q2=q2, k2=k2, drop=drawn)
    n0 = ops.LAUNCH_COUNTS["attn_keep_bits_kernel"]
    o2, l2 = ops.attn_fwd(q1, k1, v, scale, H, q2=q2, k2=k2, drop=cached)
    assert ops.LAUNCH_COUNTS["attn_keep_bits_kernel"] == n0 + 1
    assert torch.equal(o1, o2) and torch.equal(l1, l2)
    do = _rnd(B, Lq, D, seed=6)
    g1 = ops.attn_bwd(q1, k1, v, o1, l1, do, scale, H, q2=q2, k2=k2, drop=drawn)
    g2 = ops.attn_bwd(q1, k1, v, o2, l2, do, scale, H, q2=q2, k2=k2, drop=cached)
    assert ops.LAUNCH_COUNTS["attn_keep_bits_kernel"] == n0 + 1  # the backward found the forward's bits
    for a, b in zip(g1, g2):
        assert (a is None) == (b is None)
        if a is not None:
            assert torch.equal(a, b)
    # the bits themselves: keep rate, and zero weight on dropped entries is visible in the forward output changing
    bits = snap._fod_keep_bits[(11, float(P), B, H, Lq, Lk)]
    assert bits.shape == (B * H * Lq, 16 * ((Lk + 127) // 128))
    valid = torch.zeros(bits.shape[1] * 8, dtype=torch.bool, device=_dev())
    valid[:Lk] = True
    unpacked = ((bits.unsqueeze(-1) >> torch.arange(8, device=_dev(), dtype=torch.uint8)) & 1).flatten(1).bool()
    rate = float(unpacked[:, valid].float().mean())
    assert abs(rate - (1 - P)) < 5 * (P * (1 - P) / unpacked[:, valid].numel()) ** 0.5 + 2e-5, rate
