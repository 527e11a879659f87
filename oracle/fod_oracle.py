"""CPU oracle for the fod_b200 hot path — TEST INFRASTRUCTURE ONLY.

A plain-PyTorch fp32 functional restatement of the reference's forward path
(`SpatioTemporalDETR.forward` as built by `runs/_model.py:build_model`), driven by a reference
`state_dict`.  It exists so that parity tests and `bench.py`'s CPU baseline can run on the GPU box,
where `/root/reference` does not exist.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
cpu_baseline / `--impl reference` legs may import it; the product package never does.

Parity pin: the reference has no tests or golden vectors of its own, so this oracle is
pinned against outputs of the unmodified reference run in the build container
(`tests/golden/make_golden.py` -> `tests/golden/g*.pt`, `make_golden_temporal.py` -> `t*.pt` for the temporal code and
the joint-encoder mode; `tests/test_oracle_vs_golden.py`).

Every function cites the reference lines (relative to /root/reference) it restates.
Layouts follow the reference: sequences are (tokens, batch, channels).
"""
import math

import torch
import torch.nn.functional as F

from . import lsap as _lsap

D_MODEL = 256
N_HEADS = 8
LN_EPS = 1e-5


# ----------------------------------------------------------------------------- small helpers
def _lin(x, sd, name):
    return F.linear(x, sd[name + ".weight"], sd.get(name + ".bias"))


def _ln(x, sd, name):
    return F.layer_norm(x, (x.shape[-1],), sd[name + ".weight"], sd[name + ".bias"], LN_EPS)


def _mlp(x, sd, name, n_layers):
    """future_od/models/transformer.py:18-32 — Linear/ReLU stack, no activation after the last layer."""
    for i in range(n_layers):
        x = _lin(x, sd, f"{name}.layers.{i}")
        if i + 1 < n_layers:
            x = F.relu(x)
    return x


def inverse_sigmoid(x, eps=1e-5):
    """ConditionalDETR/util/misc.py:477-481."""
    x = x.clamp(0, 1)
    return torch.log(x.clamp(min=eps) / (1 - x).clamp(min=eps))


# ----------------------------------------------------------------------------- positional encodings
def _sincos_interleave(angles):
    """sin on even feature slots, cos on odd ones (paper.py:75-80, transformer.py:44-45)."""
    return torch.stack((angles[..., 0::2].sin(), angles[..., 1::2].cos()), dim=-1).flatten(-2)


def image_pos_encoding(h, w, c=D_MODEL, temperature=10000.0):
    """paper.py:57-64,75-80 on an all-ones mask: returns (c, h, w); first c/2 channels encode y."""
    half = c // 2
    ones = torch.ones((h, w), dtype=torch.float32)
    ys = ones.cumsum(0, dtype=torch.float32)
    xs = ones.cumsum(1, dtype=torch.float32)
    ys = ys / (ys[-1:, :] + 1e-6) * (2 * math.pi)
    xs = xs / (xs[:, -1:] + 1e-6) * (2 * math.pi)
    idx = torch.arange(half, dtype=torch.float32)
    dim_t = temperature ** (2 * torch.div(idx, 2, rounding_mode="floor") / half)
    ey = _sincos_interleave(ys[..., None] / dim_t)
    ex = _sincos_interleave(xs[..., None] / dim_t)
    return torch.cat((ey, ex), dim=-1).permute(2, 0, 1).contiguous()


def temporal_encoding(offsets, b, t, h, w, c=D_MODEL, temperature=10000.0, extra=0.0):
    """paper.py:66-73 (live only with PositionalEncoder(no_temporal=False); runs/_model.py:70-72 builds it off).

    offsets (B, T) of the frames the model sees, or None -> frame counter 1..T.  The code is normalised by the LAST
    frame's value (+1e-6), scaled to 2*pi, and sine-coded over all c channels.  Returns (B, T, c, h, w)."""
    mask = torch.ones((b, t, h, w), dtype=torch.float32)
    if offsets is not None:
        te = mask * offsets.float()[..., None, None] + extra
    else:
        te = mask.cumsum(1, dtype=torch.float32)
    te = te / (te[:, -1:, :, :] + 1e-6) * (2 * math.pi)
    idx = torch.arange(c, dtype=torch.float32)
    dim_t = temperature ** (2 * torch.div(idx, 2, rounding_mode="floor") / c)
    return _sincos_interleave(te[..., None] / dim_t).permute(0, 1, 4, 2, 3).contiguous()


def query_sine_embedding(ref_xy, d=D_MODEL):
    """transformer.py:35-48 — ref_xy (M, B, 2) in [0,1]; output (M, B, d) = [sine(y) ‖ sine(x)]."""
    half = d // 2
    idx = torch.arange(half, dtype=torch.float32, device=ref_xy.device)
    dim_t = 10000 ** (2 * torch.div(idx, 2, rounding_mode="floor") / half)
    ax = ref_xy[..., 0, None] * (2 * math.pi) / dim_t
    ay = ref_xy[..., 1, None] * (2 * math.pi) / dim_t
    return torch.cat((_sincos_interleave(ay), _sincos_interleave(ax)), dim=-1)


# ----------------------------------------------------------------------------- attention cores
def mha_core(q, k, v, n_heads, out_w, out_b, need_weights=False):
    """ConditionalDETR/models/attention.py:266-380 with no masks and dropout off.

    q (Lq,B,E), k (Lk,B,E), v (Lk,B,Ev): no input projections; scale = (E/heads)^-0.5 applied to q.
    Returns out (Lq,B,Ev) after out_proj and, optionally, head-averaged weights (B,Lq,Lk)."""
    lq, b, e = q.shape
    lk, ev = k.shape[0], v.shape[2]
    dh, dv = e // n_heads, ev // n_heads
    qh = (q * dh ** -0.5).reshape(lq, b * n_heads, dh).transpose(0, 1)
    kh = k.reshape(lk, b * n_heads, dh).transpose(0, 1)
    vh = v.reshape(lk, b * n_heads, dv).transpose(0, 1)
    w = torch.softmax(torch.bmm(qh, kh.transpose(1, 2)), dim=-1)
    o = torch.bmm(w, vh).transpose(0, 1).reshape(lq, b, ev)
    o = F.linear(o, out_w, out_b)
    return (o, w.view(b, n_heads, lq, lk).sum(1) / n_heads) if need_weights else (o, None)


def encoder_self_attention(x, pos, sd, pre):
    """transformer.py:401-419,457-462 + torch.nn.MultiheadAttention: q=k=x+pos, v=x, packed in-proj."""
    w, bias = sd[pre + ".attn.in_proj_weight"], sd[pre + ".attn.in_proj_bias"]
    d = x.shape[-1]
    qk_in = x + pos
    q = F.linear(qk_in, w[:d], bias[:d])
    k = F.linear(qk_in, w[d:2 * d], bias[d:2 * d])
    v = F.linear(x, w[2 * d:], bias[2 * d:])
    a, _ = mha_core(q, k, v, N_HEADS, sd[pre + ".attn.out_proj.weight"], sd[pre + ".attn.out_proj.bias"])
    x = _ln(x + a, sd, pre + ".norm1")
    hid = F.relu(_lin(x, sd, pre + ".mlp.0"))
    return _ln(x + _lin(hid, sd, pre + ".mlp.3"), sd, pre + ".norm2")


def egodeep_attention(x, pos, ego, sd, pre):
    """transformer.py:85-119 — attention of every token over the IMU key(s) `ego` (K, B, D)."""
    q = _lin(x, sd, pre + ".query_content") + _lin(pos, sd, pre + ".query_pos")
    out, _ = mha_core(q, _lin(ego, sd, pre + ".key"), _lin(ego, sd, pre + ".value"), N_HEADS,
                      sd[pre + ".fun.out_proj.weight"], sd[pre + ".fun.out_proj.bias"])
    if (pre + ".norm1.weight") in sd:  # Dff given -> use_mlp
        out = _ln(out + out, sd, pre + ".norm1")  # eval-mode dropout is the identity: out + out
        hid = F.relu(_lin(out, sd, pre + ".mlp.0"))
        out = _ln(out + _lin(hid, sd, pre + ".mlp.3"), sd, pre + ".norm2")
    return out


def encoder(x, pos, ego, sd, pre, n_layers=6):
    """transformer.py:449-512 — x,pos (S, N, D); ego (K, N, D) or None."""
    for i in range(n_layers):
        lp = f"{pre}.layers.{i}"
        x = encoder_self_attention(x, pos, sd, lp + ".self_attn")
        if ego is not None:
            x = _ln(x + egodeep_attention(x, pos, ego, sd, lp + ".egodeep_attend"), sd, lp + ".norm_eda")
    return x


# ----------------------------------------------------------------------------- decoder
def slot_to_slot(c, p, sd, pre):
    """transformer.py:61-82 — decoder self-attention; content and positional parts projected separately."""
    q = _lin(c, sd, pre + ".query_content") + _lin(p, sd, pre + ".query_pos")
    k = _lin(c, sd, pre + ".key_content") + _lin(p, sd, pre + ".key_pos")
    v = _lin(c, sd, pre + ".value")
    return mha_core(q, k, v, N_HEADS, sd[pre + ".fun.out_proj.weight"], sd[pre + ".fun.out_proj.bias"])[0]


def slot_to_image(c, qpos, qsine, img, img_pos, is_first, sd, pre, need_weights=False):
    """transformer.py:122-181 — conditional cross-attention: per head q=[content ‖ sine], k=[content ‖ pos]."""
    m, b, d = c.shape
    n = img.shape[0]
    dh = d // N_HEADS
    v = _lin(img, sd, pre + ".value")
    qc = _lin(c, sd, pre + ".query_content")
    if is_first:
        qc = qc + _lin(qpos, sd, pre + ".query_pos")
    qs = _lin(qsine, sd, pre + ".query_sine")
    q = torch.cat((qc.view(m, b, N_HEADS, dh), qs.view(m, b, N_HEADS, dh)), dim=3).reshape(m, b, 2 * d)
    kp = _lin(img_pos, sd, pre + ".key_pos")
    kc = _lin(img, sd, pre + ".key_content")
    if is_first:
        kc = kc + kp
    k = torch.cat((kc.view(n, b, N_HEADS, dh), kp.view(n, b, N_HEADS, dh)), dim=3).reshape(n, b, 2 * d)
    return mha_core(q, k, v, N_HEADS, sd[pre + ".fun.out_proj.weight"], sd[pre + ".fun.out_proj.bias"],
                    need_weights=need_weights)


def decoder_layer(c, qpos, qsine, images, poses, is_first, sd, pre, attn_sink=None):
    """transformer.py:242-312 (use_slotstates=False, use_egodeep=False as built by runs/_model.py:48-55)."""
    c = _ln(c + slot_to_slot(c, qpos, sd, pre + ".self_attend"), sd, pre + ".norm_sa")
    for i, (img, ipos) in enumerate(zip(images, poses)):
        upd, wts = slot_to_image(c, qpos, qsine, img, ipos, is_first, sd, f"{pre}.image_attend.{i}",
                                 need_weights=attn_sink is not None)
        if attn_sink is not None:
            attn_sink.append(wts)
        c = _ln(c + upd, sd, f"{pre}.norm_ia.{i}")
    hid = F.relu(_lin(c, sd, pre + ".feedforward.0"))
    return _ln(c + _lin(hid, sd, pre + ".feedforward.3"), sd, pre + ".norm_out")


def decoder(qpos, images, poses, sd, pre, n_layers=6, first_layer_special=True, attn_sink=None):
    """transformer.py:332-398 — returns hs (layers, B, M, D) and reference points (B, M, 2)."""
    ref = _mlp(qpos, sd, pre + ".ref_point_head", 2).sigmoid()  # (M, B, 2)
    sine0 = query_sine_embedding(ref)
    c = torch.zeros_like(qpos)
    levels = []
    for i in range(n_layers):
        special = i == 0 and first_layer_special
        qsine = sine0 if special else _mlp(c, sd, pre + ".query_scale", 2) * sine0
        c = decoder_layer(c, qpos, qsine, images, poses, special, sd, f"{pre}.layers.{i}", attn_sink)
        levels.append(_ln(c, sd, pre + ".norm"))
    return torch.stack(levels).transpose(1, 2), ref.transpose(0, 1)


def detect(frames, poses, sd, pre, n_layers=6):
    """paper.py:352-419 — frames/poses: lists of (S, B, D), current frame first."""
    b = frames[0].shape[1]
    qpos = sd[pre + ".query_embed.weight"].unsqueeze(1).repeat(1, b, 1)
    hs, ref = decoder(qpos, frames, poses, sd, pre + ".decoder", n_layers)
    ref_logit = inverse_sigmoid(ref)
    boxes = []
    for lvl in range(hs.shape[0]):
        t = _mlp(hs[lvl], sd, pre + ".bbox_embed", 3).clone()
        t[..., :2] = t[..., :2] + ref_logit
        boxes.append(t.sigmoid())
    boxes = torch.stack(boxes)
    logits = _lin(hs, sd, pre + ".class_embed")
    out = {"pred_logits": logits[-1], "pred_boxes": boxes[-1], "hs": hs,
           "aux_outputs": [{"pred_logits": a, "pred_boxes": c} for a, c in zip(logits[:-1], boxes[:-1])]}
    return out


def detector(features, pos, sd, pre, num_images=2, all_passes=False):
    """paper.py:331-350 ("attend one at a time"): frame l attends itself and the num_images-1 previous ones.

    Only the last pass reaches the output; all_passes=True also runs the dead ones, as
    the reference does, for timing purposes."""
    _, l = features.shape[:2]
    toks = features.flatten(3).permute(1, 3, 0, 2)  # (L, HW, B, C)
    ptok = pos.flatten(3).permute(1, 3, 0, 2)
    out = None
    for cur in (range(l) if all_passes else [l - 1]):
        lst = [toks[cur - j] for j in range(num_images) if cur - j >= 0]
        out = detect(lst, [ptok[cur] for _ in lst], sd, pre)
    return out


# ----------------------------------------------------------------------------- backbone (library code)
class _FrozenBN(torch.nn.Module):
    """ConditionalDETR/models/backbone.py:27-63 — affine with fixed statistics, eps 1e-5."""

    def __init__(self, n):
        super().__init__()
        for name, val in (("weight", torch.ones(n)), ("bias", torch.zeros(n)),
                          ("running_mean", torch.zeros(n)), ("running_var", torch.ones(n))):
            self.register_buffer(name, val)

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        state_dict.pop(prefix + "num_batches_tracked", None)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def forward(self, x):
        scale = self.weight * (self.running_var + 1e-5).rsqrt()
        shift = self.bias - self.running_mean * scale
        return x * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)


_BACKBONES = {}


def backbone_features(images, sd, pre):
    """paper.py:83-116 — torchvision ResNet-50 body (FrozenBN) up to layer4, then the 1x1 input_proj."""
    import torchvision

    key = id(sd)
    if key not in _BACKBONES:
        net = torchvision.models.resnet50(weights=None, norm_layer=_FrozenBN)
        del net.fc, net.avgpool
        body = {k[len(pre + ".body."):]: v for k, v in sd.items() if k.startswith(pre + ".body.")}
        net.load_state_dict(body, strict=True)
        _BACKBONES.clear()
        _BACKBONES[key] = net.eval()
    net = _BACKBONES[key]
    x = net.maxpool(net.relu(net.bn1(net.conv1(images))))
    x = net.layer4(net.layer3(net.layer2(net.layer1(x))))
    return F.conv2d(x, sd[pre + ".input_proj.weight"], sd[pre + ".input_proj.bias"])


def core_forward(images, imu, sd, pre="_model", all_passes=False, temporal=None, joint=False):
    """paper.py:448-485 + :133-170 — images (B, L, 3, H, W), imu (B, L, 14) or None.

    temporal=None is PositionalEncoder(no_temporal=True) (what runs/_model.py builds); temporal={"offsets": (B, L)
    tensor or None, "extra": float} adds the temporal sine code to the encoding handed to the joint encoder and the
    decoder (paper.py:50-55,473-482).  In the per-frame configuration that code is constant over the keys of every
    softmax it enters (each cross-attention sees ONE frame's encoding, paper.py:381-382), so it cancels; it is live
    only with joint=True — the optional JointEncoder (paper.py:180-203) run with the per-frame encoder's layers."""
    images = images[:, :-1]
    b, t = images.shape[:2]
    feats = backbone_features(images.flatten(0, 1), sd, pre + ".separate_encoder.backbone")
    _, c, h, w = feats.shape
    pos = image_pos_encoding(h, w, c)
    ego = None
    if imu is not None:
        e = imu[:, :-1]
        e = F.linear(F.relu(_lin(e, sd, pre + ".separate_encoder.imu_layers.0")),
                     sd[pre + ".separate_encoder.imu_layers.2.weight"], sd[pre + ".separate_encoder.imu_layers.2.bias"])
        ego = e.reshape(1, b * t, c)
    x = feats.flatten(2).permute(2, 0, 1)  # (HW, BT, C)
    p = pos.flatten(1).t().unsqueeze(1).expand(-1, b * t, -1)
    x = encoder(x, p, ego, sd, pre + ".separate_encoder.transformer")
    feats = x.permute(1, 2, 0).reshape(b, t, c, h, w)
    pos_bt = pos.view(1, 1, c, h, w).expand(b, t, -1, -1, -1)
    if temporal is not None:
        offs = temporal.get("offsets")
        pos_bt = pos_bt + temporal_encoding(None if offs is None else offs[:, :-1], b, t, h, w, c,
                                            extra=temporal.get("extra", 0.0))
    if joint:
        xj = feats.permute(3, 4, 1, 0, 2).reshape(h * w * t, b, c)          # "b l c h w -> (h w l) b c"
        pj = pos_bt.permute(3, 4, 1, 0, 2).reshape(h * w * t, b, c)
        ej = ego.view(b, t, c).permute(1, 0, 2) if ego is not None else None  # "b l c -> l b c": T keys per clip
        xj = encoder(xj, pj, ej, sd, pre + ".separate_encoder.transformer")
        feats = xj.view(h, w, t, b, c).permute(3, 2, 4, 0, 1)
    out = detector(feats, pos_bt, sd, pre + ".detector", all_passes=all_passes)
    out["encoder_features"] = feats
    return out


# ----------------------------------------------------------------------------- set-prediction loss
def cxcywh_to_xyxy(b):
    """ConditionalDETR/util/box_ops.py:17-21."""
    cx, cy, w, h = b.unbind(-1)
    return torch.stack((cx - 0.5 * w, cy - 0.5 * h, cx + 0.5 * w, cy + 0.5 * h), dim=-1)


def pairwise_giou(a, b):
    """ConditionalDETR/util/box_ops.py:32-69 — a (N,4), b (M,4) xyxy -> (N,M)."""
    assert (a[:, 2:] >= a[:, :2]).all() and (b[:, 2:] >= b[:, :2]).all()
    area_a = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
    area_b = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    wh = (torch.min(a[:, None, 2:], b[:, 2:]) - torch.max(a[:, None, :2], b[:, :2])).clamp(min=0)
    inter = wh[..., 0] * wh[..., 1]
    union = area_a[:, None] + area_b - inter
    iou = inter / union
    hull = (torch.max(a[:, None, 2:], b[:, 2:]) - torch.min(a[:, None, :2], b[:, :2])).clamp(min=0)
    hull_area = hull[..., 0] * hull[..., 1]
    return iou - (hull_area - union) / hull_area


def matcher_cost(logits, boxes, tgt_labels, tgt_boxes, w_class=2.0, w_bbox=5.0, w_giou=2.0):
    """ConditionalDETR/models/matcher.py:58-82 for ONE sample: logits (M,C), boxes (M,4) -> (M,Nt) fp32."""
    prob = logits.sigmoid()
    alpha, gamma = 0.25, 2.0
    neg = (1 - alpha) * (prob ** gamma) * (-(1 - prob + 1e-8).log())
    pos = alpha * ((1 - prob) ** gamma) * (-(prob + 1e-8).log())
    c_class = pos[:, tgt_labels] - neg[:, tgt_labels]
    c_bbox = torch.cdist(boxes, tgt_boxes, p=1)
    c_giou = -pairwise_giou(cxcywh_to_xyxy(boxes), cxcywh_to_xyxy(tgt_boxes))
    return w_bbox * c_bbox + w_class * c_class + w_giou * c_giou


def hungarian_match(logits, boxes, targets):
    """ConditionalDETR/models/matcher.py:40-87 — per-sample (pred_idx ascending, tgt_idx) int64 pairs.

    The reference builds the cross-batch cost and keeps the block diagonal; per-sample blocks are the same
    numbers because every op is element-wise in (query, target)."""
    out = []
    for b, t in enumerate(targets):
        if t["labels"].numel() == 0:
            out.append((torch.zeros(0, dtype=torch.int64), torch.zeros(0, dtype=torch.int64)))
            continue
        cost = matcher_cost(logits[b].float(), boxes[b].float(), t["labels"], t["boxes"])
        r, c = _lsap.linear_sum_assignment(cost.double().cpu().numpy())  # matcher.py:84: C.cpu()
        out.append((torch.as_tensor(r, dtype=torch.int64), torch.as_tensor(c, dtype=torch.int64)))
    return out


def to_detr_targets(h, w, active, boxes, classes):
    """st_detr.py:237-263 — dense xyxy pixel slots -> per-sample normalised cxcywh lists."""
    cxcywh = torch.cat((0.5 * (boxes[..., 0:2] + boxes[..., 2:4]), boxes[..., 2:4] - boxes[..., 0:2]), dim=2)
    cxcywh = cxcywh * torch.tensor([1 / w, 1 / h, 1 / w, 1 / h]).view(1, 1, 4)
    return [{"labels": c[a == 1], "boxes": bx[a == 1]} for c, bx, a in zip(classes, cxcywh, active)]


def focal_loss(logits, onehot, num_boxes, alpha=0.25, gamma=2.0):
    """ConditionalDETR/models/segmentation.py:198-223."""
    p = logits.sigmoid()
    ce = F.binary_cross_entropy_with_logits(logits, onehot, reduction="none")
    p_t = p * onehot + (1 - p) * (1 - onehot)
    loss = ce * (1 - p_t) ** gamma
    loss = (alpha * onehot + (1 - alpha) * (1 - onehot)) * loss
    return loss.mean(1).sum() / num_boxes


def level_losses(logits, boxes, targets, indices, num_boxes, num_classes, log):
    """set_criterion.py:36-115 for one decoder level."""
    bsz, m, _ = logits.shape
    bi = torch.cat([torch.full_like(s, i) for i, (s, _) in enumerate(indices)])
    si = torch.cat([s for s, _ in indices])
    tcls = torch.cat([t["labels"][j] for t, (_, j) in zip(targets, indices)])
    tbox = torch.cat([t["boxes"][j] for t, (_, j) in zip(targets, indices)], dim=0)
    onehot = torch.zeros(bsz, m, num_classes)
    onehot[bi, si, tcls] = 1.0
    out = {"loss_ce": focal_loss(logits, onehot, num_boxes) * m}
    if log:
        matched = logits[bi, si]
        if tcls.numel() == 0:
            acc = torch.zeros([])
        else:
            acc = (matched.argmax(1) == tcls).float().sum() * (100.0 / tcls.numel())
        out["class_error"] = 100 - acc
    n_tgt = torch.tensor([len(t["labels"]) for t in targets], dtype=torch.float32)
    card = (logits.max(-1)[0] > 0.5).sum(1).float()
    out["cardinality_error"] = F.l1_loss(card, n_tgt)
    sbox = boxes[bi, si]
    out["loss_bbox"] = (sbox - tbox).abs().sum() / num_boxes
    giou = torch.diag(pairwise_giou(cxcywh_to_xyxy(sbox), cxcywh_to_xyxy(tbox))) if sbox.numel() else sbox.new_zeros(0)
    out["loss_giou"] = (1 - giou).sum() / num_boxes
    return out


def criterion(out, targets, num_classes=8, world_num_boxes=None, weights=(2.0, 5.0, 2.0)):
    """set_criterion.py:172-217 ("per level" matching) + st_detr.py:169-188 weighting.

    Returns (loss, stats dict, per-level assignment lists)."""
    n = float(sum(len(t["labels"]) for t in targets)) if world_num_boxes is None else float(world_num_boxes)
    num_boxes = max(n, 1.0)
    levels = [(out["pred_logits"], out["pred_boxes"])] + [(a["pred_logits"], a["pred_boxes"]) for a in out["aux_outputs"]]
    w_ce, w_bbox, w_giou = weights
    total = 0.0
    stats, all_idx = {}, []
    for i, (lg, bx) in enumerate(levels):
        idx = hungarian_match(lg.detach(), bx.detach(), targets)
        all_idx.append(idx)
        ls = level_losses(lg, bx, targets, idx, num_boxes, num_classes, log=(i == 0))
        total = total + w_ce * ls["loss_ce"] + w_bbox * ls["loss_bbox"] + w_giou * ls["loss_giou"]
        if i == 0:
            stats = {"labels": ls["loss_ce"] * w_ce, "box_l1": ls["loss_bbox"] * w_bbox,
                     "box_giou": ls["loss_giou"] * w_giou, "cardinality": ls["cardinality_error"],
                     "class_error": ls["class_error"]}
    return total, stats, all_idx


# ----------------------------------------------------------------------------- post-processing / AP bookkeeping
_SIZE_DELIMS = ((1 / 24) * (1 / 64), (1 / 4) * (1 / 12))  # od_map.py:20-23, fractions of H*W


def _size_onehot(boxes, imsize):
    """od_map.py:163-181 — (B, K, 4) xyxy pixels -> bool (B, K, 4) [all, small, medium, large]."""
    h, w = imsize
    area = (boxes[..., 2] - boxes[..., 0]) * (boxes[..., 3] - boxes[..., 1])
    lo, hi = _SIZE_DELIMS[0] * h * w, _SIZE_DELIMS[1] * h * w
    return torch.stack((torch.ones_like(area, dtype=torch.bool), area <= lo, (lo < area) & (area <= hi), hi < area), dim=2)


def od_map_stuffs(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk=50):
    """od_map.py:214-287 — greedy per-class matching of the top-50 detections at 10 IoU thresholds.

    pred_boxes (B,M,4) xyxy px, scores (B,M,C) with the generic class last, dense annotation slots (B,N',...)."""
    bsz, _, ncls = scores.shape
    thr = torch.arange(0.50, 1.00, 0.05)
    keep = anno_active.any(dim=0) == 1  # od_map.py:40-45: drop slots unused by every sample, keep slot 0
    keep[0] = True
    ab, ac, aa = anno_boxes[:, keep], anno_classes[:, keep], anno_active[:, keep]
    n = ab.shape[1]
    # od_map.py:48-71 IoU with +1e-7 smoothing
    p, a = pred_boxes[:, :, None, :], ab[:, None, :, :]
    area_p = F.relu(p[..., 2] - p[..., 0]) * F.relu(p[..., 3] - p[..., 1])
    area_a = F.relu(a[..., 2] - a[..., 0]) * F.relu(a[..., 3] - a[..., 1])
    inter = F.relu(torch.min(p[..., 2], a[..., 2]) - torch.max(p[..., 0], a[..., 0])) * \
        F.relu(torch.min(p[..., 3], a[..., 3]) - torch.max(p[..., 1], a[..., 1]))
    iou = (inter + 1e-7) / (area_p + area_a - inter + 1e-7)  # (B, M, N)
    conf, order = scores.sort(dim=1, descending=True)  # od_map.py:88-105
    conf, order = conf[:, :topk], order[:, :topk]      # (B, K, C)
    k = conf.shape[1]
    cls_ok = torch.cat((ac[:, None, :] == torch.arange(ncls - 1)[None, :, None],
                        torch.ones((bsz, 1, n), dtype=torch.bool)), dim=1)
    avail = (aa[:, None, :] == 1) & cls_ok  # (B, C, N)  od_map.py:108-132
    iou_k = iou[:, :, None, :].expand(-1, -1, ncls, -1).gather(1, order[..., None].expand(-1, -1, -1, n))
    iou_k = torch.where(avail[:, None], iou_k, torch.zeros_like(iou_k))  # (B, K, C, N)
    alive = torch.ones((bsz, len(thr), ncls, n))
    positive = torch.zeros((bsz, len(thr), k, ncls), dtype=torch.bool)
    for m in range(k):  # od_map.py:267-277 — a claimed annotation is zeroed for all later detections
        best, arg = (iou_k[:, None, m] * alive).max(dim=3)
        hit = best >= thr[None, :, None]
        positive[:, :, m] = hit
        cur = alive.gather(3, arg[..., None])
        alive.scatter_(3, arg[..., None], torch.where(hit[..., None], torch.zeros_like(cur), cur))
    confs = conf.reshape(bsz * k, ncls).t().reshape(1, ncls, bsz * k).repeat(len(thr), 1, 1)
    is_pos = positive.permute(1, 3, 0, 2).reshape(len(thr), ncls, bsz * k)
    sz = _size_onehot(pred_boxes, imsize)  # (B, M, 4)
    sz = sz[:, :, None, :].expand(-1, -1, ncls, -1).gather(1, order[..., None].expand(-1, -1, -1, 4))
    size_cat = sz.reshape(bsz * k, ncls, 4).permute(1, 2, 0)
    num_annos = (avail[..., None] & _size_onehot(ab, imsize)[:, None]).sum(dim=(0, 2))  # (C, 4) od_map.py:199-211
    return confs, is_pos, size_cat, num_annos


def average_precision(confs, is_positive, size_categories, num_annos):
    """od_map.py:288-312 (_get_ap) for one IoU threshold — confs, is_positive (C, N); size_categories (C, S, N) bool;
    num_annos (C, S, iterations).  AP = sum over positives of precision-at-rank / number of annotations; 0/0 -> NaN and
    x/0 -> inf are kept (the caller nan-means over them)."""
    c, s, n = size_categories.shape
    order = confs.argsort(dim=1, descending=True).view(c, 1, n).expand(-1, s, -1)
    pos = (is_positive.view(c, 1, n) * size_categories).gather(2, order)
    cat = size_categories.gather(2, order)
    precision = pos.cumsum(dim=2) / (cat.cumsum(dim=2) + 1e-5)
    return (precision * pos).sum(dim=2) / num_annos.sum(dim=2)


def aggregate_ap(confs, is_positive, size_categories, num_annos):
    """od_map.py:315-364 — confs, is_positive (T, C, N) stacked over iterations; returns the reference's dict
    (generic class last; nan-means over classes / thresholds) as tensors."""
    import numpy as np
    with torch.no_grad():
        ap = torch.stack([average_precision(confs[t], is_positive[t], size_categories, num_annos)
                          for t in range(confs.shape[0])])
    a = ap.numpy()
    with np.errstate(all="ignore"):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", RuntimeWarning)
            out = {"all": a[:, :-1], "classavg": np.nanmean(a[:, :-1], axis=1), "threshavg": np.nanmean(a[:, :-1], axis=0),
                   "classavg threshavg": np.nanmean(a[:, :-1], axis=(0, 1)), "generic": a[:, -1],
                   "generic threshavg": np.nanmean(a[:, -1], axis=0)}
    return {k: torch.as_tensor(np.ascontiguousarray(v)) for k, v in out.items()}


def post_process(pred_logits, pred_boxes, data):
    """st_detr.py:190-234 for the single-output case (pred_* already (B, M, .))."""
    h, w = data["video"].shape[-2:]
    sc = pred_logits.sigmoid()
    sc = torch.cat((sc, sc.max(dim=2, keepdim=True)[0]), dim=2)
    bx = pred_boxes * torch.tensor([w, h, w, h]).view(1, 1, 4)
    bx = torch.cat((bx[..., 0:2] - 0.5 * bx[..., 2:4], bx[..., 0:2] + 0.5 * bx[..., 2:4]), dim=-1)
    stuffs = od_map_stuffs(bx, sc, data["boxes"], data["classes"], data["active"], (h, w))
    return stuffs, {"class_scores": sc[:, None, None], "boxes": bx[:, None, None]}


IMU_KEYS = ("translation", "acceleration", "rotation", "rotation_rate", "speed")


def forward(data, sd, all_passes=False, with_od_map=True, temporal=None):
    """st_detr.py:98-167 — returns (outputs, None, loss, stats, od_map_stuffs) plus the raw core dict."""
    video = data["video"]
    imu = torch.cat([data[k] for k in IMU_KEYS], dim=2) if data.get("translation") is not None else None
    core = core_forward(video, imu, sd, all_passes=all_passes, temporal=temporal)
    h, w = video.shape[-2:]
    targets = to_detr_targets(h, w, data["active"], data["boxes"], data["classes"])
    loss, stats, idx = criterion(core, targets)
    stuffs, outputs = post_process(core["pred_logits"], core["pred_boxes"], data) if with_od_map else (None, {})
    outputs["moods"] = [["model happy" for _ in range(video.shape[1])] for _ in range(video.shape[0])]
    core["indices"] = idx
    return outputs, None, loss, stats, stuffs, core
