"""Top of the drop-in: SeparateEncoder / FuturePredCore (paper.py:119-170, 432-485), the on-device set criterion
(set_criterion.py + ConditionalDETR matcher), post-processing (st_detr.py:190-234, utils/od_map.py:214-287) and
SpatioTemporalDETR with the reference's forward signature and return tuple (st_detr.py:98-167), plus
`build_model(args, detr_args)` mirroring runs/_model.py:14-83.
"""
import os
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import autograd as A
from . import lib, ops
from . import model as M
from . import detector as D
from .dist import global_num_boxes, is_dist

BF16 = torch.bfloat16


@dataclass
class SpatioTemporalDETRArgs:
    """Field-for-field mirror of future_od/models/st_detr.py:12-55."""
    num_classes: int
    masks: bool = False
    lr_backbone: float = 1e-5
    lr: float = 1e-4
    weight_decay: float = 1e-4
    max_norm: float = 0.1
    backbone: str = "resnet50"
    dilation: bool = False
    position_embedding: str = "sine"
    pretrained_backbone: bool = True
    enc_layers: int = 6
    dec_layers: int = 6
    dim_feedforward: int = 2048
    hidden_dim: int = 256
    dropout: float = 0.1
    enc_nheads: int = 8
    nheads: int = 8
    num_queries: int = 300
    pre_norm: bool = False
    set_cost_class: float = 2.0
    set_cost_bbox: float = 5.0
    set_cost_giou: float = 2.0
    aux_loss: bool = True
    cls_loss_coef: float = 2.0
    bbox_loss_coef: float = 5.0
    giou_loss_coef: float = 2.0
    focal_alpha: float = 0.25
    no_imu_speed: bool = False
    encode_offset: bool = False


class SeparateEncoder(nn.Module):
    """paper.py:119-170 — per-frame backbone + IMU MLP + 6-layer encoder on (B*T) sequences of h*w tokens."""

    def __init__(self, backbone, imu_layers=None, transformer=None):
        super().__init__()
        self.backbone = backbone
        self.imu_layers = imu_layers
        self.transformer = transformer

    def _imu(self, imu):
        l0, l2 = self.imu_layers[0], self.imu_layers[2]
        k = l0.weight.shape[1]
        kp = (k + 7) // 8 * 8  # TMA rows need 16-byte strides: pad 14 -> 16 input features with zeros
        x = F.pad(imu.reshape(-1, k).float(), (0, kp - k)).to(BF16)
        w0 = F.pad(l0.weight, (0, kp - k))
        h = A.linear(x, w0, l0.bias, M._q(w0), None, 0, True, False)
        return M._lin(h, l2)

    def forward_tokens(self, images, pos_encoder, imu=None):
        """images (B,T,3,H,W) -> encoded tokens (B*T*S, D) bf16 in (b, t, h, w) row order, pos (S, D) bf16, (h, w)."""
        b, t = images.shape[:2]
        bk = self.backbone
        ih, iw = images.shape[3:]
        if (images.is_cuda and images.dtype == torch.float32 and not images.requires_grad and images.shape[2] == 3
                and getattr(bk, "fold_frozen_bn", False) and getattr(bk, "compute_dtype", None) == BF16 and iw % 8 == 0
                and images.stride()[2:] == (ih * iw, iw, 1) and images.stride(0) % 4 == 0 and images.stride(1) % 4 == 0
                and (ih * iw) % 4 == 0 and images.data_ptr() % 16 == 0):
            # (B, T, 3, H, W) fp32 planar (possibly a frame slice of the loader's tensor) -> channels-last bf16 in one pass
            # (the padded 4-channel form feeds the TMA stem convolution; odd heights keep the 3-channel form)
            frames = ops.frames_to_nhwc4p_bf16(images) if (ih % 2 == 0 and _STEM_P4) else ops.frames_to_nhwc_bf16(images)
        else:
            frames = images.flatten(0, 1)
        x, h, w = self.backbone.forward_tokens(frames)
        d = x.shape[1]
        pos_bf = pos_encoder.spatial_tokens(h, w, d, x.device).to(BF16)
        ego = None
        if imu is not None and self.imu_layers is not None:
            ego = self._imu(imu)  # (B*T, D): ONE key per sequence (paper.py:162)
        if self.transformer is not None:
            x = self.transformer(x, pos_bf, ego, b * t, h * w)
        return x, pos_bf, ego, h, w

    def forward(self, images, pos_encoder, imu=None, temporal_offsets=None):
        """Reference-shaped outputs: features (B,T,D,h,w) fp32 and egodeep (B,T,D) (paper.py:133-170)."""
        b, t = images.shape[:2]
        x, _, ego, h, w = self.forward_tokens(images, pos_encoder, imu)
        feats = x.view(b, t, h, w, -1).permute(0, 1, 4, 2, 3).float()
        return feats, (ego.view(b, t, -1).float() if ego is not None else None)


class FuturePredCore(nn.Module):
    """paper.py:432-485."""

    def __init__(self, separate_encoder, joint_encoder, detector, pos_encoder):
        super().__init__()
        self.separate_encoder = separate_encoder
        self.joint_encoder = joint_encoder
        self.detector = detector
        self.pos_encoder = pos_encoder

    def forward(self, images, imu=None, temporal_offsets=None):
        assert images.dim() == 5, "expected (B, L, 3, H, W)"  # paper.py:463
        ops.begin_forward(self.training and torch.is_grad_enabled())
        images = images[:, :-1]  # the last ("future") frame is never seen by the model (paper.py:463-468)
        if imu is not None:
            imu = imu[:, :-1]
        if temporal_offsets is not None:
            temporal_offsets = temporal_offsets[:, :-1]
        b, t = images.shape[:2]
        x, pos_bf, ego, h, w = self.separate_encoder.forward_tokens(images, self.pos_encoder, imu)
        s, d = h * w, x.shape[1]
        if self.joint_encoder is not None:
            # Temporal code (paper.py:50-55,66-73; off in runs/_model.py:70-72).  It is one vector per (clip, frame),
            # so it only matters where tokens of DIFFERENT frames share a softmax: the joint encoder.  In the decoder
            # every cross-attention sees one frame's encoding (paper.py:381-382); there the code shifts all keys by
            # the same vector and cancels exactly (tests/golden t1: reference ON == OFF), so it is not materialised.
            tcode = pos_f32 = None
            if not self.pos_encoder.no_temporal:
                tcode = self.pos_encoder.temporal_channels(temporal_offsets, b, t, d, x.device)
                pos_f32 = self.pos_encoder.spatial_tokens(h, w, d, x.device)
            x = self.joint_encoder(x, pos_bf, ego, b, t, s, tcode, pos_f32)
        self.last_encoder_tokens = x.detach()  # detached: a live graph would pin last step's AccumulateGrad nodes
        # (b, t, s) rows -> one contiguous (B*S, D) token matrix per frame for the decoder's K/V GEMMs
        per_frame = x.view(b, t, s, d).transpose(0, 1).contiguous()
        frames = [per_frame[l].reshape(b * s, d) for l in range(t)]
        out = self.detector(frames, pos_bf, b, s)
        ops.end_forward()
        moods = [["model happy" for _ in range(t + 1)] for _ in range(b)]
        return out, moods


class JointEncoder(nn.Module):
    """paper.py:180-203 — the same encoder layers over ALL T*h*w tokens of a clip jointly (token order (h w l)),
    IMU keys (T per clip).  Optional mode, not built by runs/_model.py; exercises attention at S = T*h*w."""

    def __init__(self, transformer):
        super().__init__()
        self.transformer = transformer

    def forward(self, x, pos_bf, ego, b, t, s, tcode=None, pos_f32=None):
        """tcode (B, T, D) fp32: optional temporal code, added to the spatial one (pos_f32, (S, D)) per (clip, frame)."""
        d = x.shape[1]
        xt = x.view(b, t, s, d).permute(0, 2, 1, 3).reshape(b * s * t, d)           # (b, hw, l) token order
        if tcode is None:
            pos_j = pos_bf.view(s, 1, d).expand(s, t, d).reshape(s * t, d).contiguous()  # spatial code repeated over l
        else:  # one row per token: spatial[hw] + temporal[b, l], summed in fp32 as the reference does
            sp = pos_f32 if pos_f32 is not None else pos_bf.float()
            pos_j = (sp.view(1, s, 1, d) + tcode.view(b, 1, t, d)).to(BF16).reshape(b * s * t, d)
        ego_j = ego.view(b, t, d) if ego is not None else None
        y = self.transformer(xt, pos_j, ego_j, b, s * t)
        return y.view(b, s, t, d).permute(0, 2, 1, 3).reshape(b * t * s, d)


# ------------------------------------------------------------------------------------------------ criterion
class HungarianMatcher(nn.Module):
    """ConditionalDETR/models/matcher.py:23-87 on the device: cost blocks + LSAP in one kernel launch (fod_match).
    Returns the reference's list of (pred_idx, tgt_idx) int64 pairs (prediction indices ascending)."""

    def __init__(self, cost_class=1.0, cost_bbox=1.0, cost_giou=1.0):
        super().__init__()
        assert cost_class != 0 or cost_bbox != 0 or cost_giou != 0, "all costs cant be 0"
        self.cost_class, self.cost_bbox, self.cost_giou = cost_class, cost_bbox, cost_giou

    @torch.no_grad()
    def forward(self, outputs, targets):
        """outputs: {"pred_logits" (B,M,C), "pred_boxes" (B,M,4)}; targets: list of {"labels" (Nt,), "boxes" (Nt,4)}."""
        logits, boxes = outputs["pred_logits"], outputs["pred_boxes"]
        bsz, m = logits.shape[:2]
        nt_max = max(1, max(len(t["labels"]) for t in targets))
        nt_max = (nt_max + 7) // 8 * 8
        tb = torch.zeros((bsz, nt_max, 4), dtype=torch.float32, device=logits.device)
        tl = torch.zeros((bsz, nt_max), dtype=torch.int32, device=logits.device)
        cnt = torch.tensor([len(t["labels"]) for t in targets], dtype=torch.int32, device=logits.device)
        for i, t in enumerate(targets):
            n = len(t["labels"])
            tb[i, :n] = t["boxes"].float()
            tl[i, :n] = t["labels"].to(torch.int32)
        r = ops.match(logits[None], boxes[None], tb, tl, cnt, self.cost_class, self.cost_bbox, self.cost_giou)
        check_match_status(r["status"])
        pi, ti = r["pred_idx"][0].cpu(), r["tgt_idx"][0].cpu()
        return [(pi[i][pi[i] >= 0], ti[i][ti[i] >= 0]) for i in range(bsz)]


def check_match_status(status):
    """Raise what the reference raises (host sync: call outside the hot loop or at the end of a step)."""
    s = int(status.item())
    if s & 1:
        raise ValueError("matrix contains invalid numeric entries")  # scipy, via matcher.py:86
    if s & 2:
        raise AssertionError("degenerate boxes (x1 < x0 or y1 < y0)")  # box_ops.py:59-60
    if s & 4:
        raise ValueError("cost matrix is infeasible")


class SetCriterion(nn.Module):
    """future_od/models/set_criterion.py:14-217, matching_mode="per level", losses labels+boxes+cardinality: all six
    levels are matched in one launch and their losses AND gradients come from one fused kernel; no host sync."""

    def __init__(self, num_classes, matcher, weight_dict, focal_alpha, losses, matching_mode="per level"):
        super().__init__()
        if matching_mode != "per level":
            raise NotImplementedError(matching_mode)
        self.num_classes, self.matcher, self.weight_dict = num_classes, matcher, weight_dict
        self.focal_alpha, self.losses = focal_alpha, losses
        self.last_status = None
        self.last_match = None

    def forward(self, all_logits, all_boxes, tgt_boxes, tgt_labels, tgt_count, total_count, distributed=False,
                num_boxes=None):
        """all_logits (L,B,M,C), all_boxes (L,B,M,4) fp32 with the main output last.  Returns (loss, stats).
        num_boxes: the already reduced global target count (SpatioTemporalDETR starts that all-reduce before the
        forward pass so that no rank waits for another in the middle of the step)."""
        if num_boxes is None:
            num_boxes = global_num_boxes(total_count, distributed)  # set_criterion.py:185-193, stays on the device
        wd, mt = self.weight_dict, self.matcher
        weights = (wd["loss_ce"], wd["loss_bbox"], wd["loss_giou"], mt.cost_class, mt.cost_bbox, mt.cost_giou,
                   self.focal_alpha)
        status = torch.zeros((1,), dtype=torch.int32, device=all_logits.device)
        out, match_tgt = A.SetLossFn.apply(all_logits, all_boxes, tgt_boxes, tgt_labels, tgt_count, num_boxes, weights,
                                           status)
        self.last_status, self.last_match = status, match_tgt
        o = out.detach()
        stats = {"labels": o[1], "box_l1": o[2], "box_giou": o[3], "cardinality": o[4], "class_error": o[5]}
        return out[0], stats


# ------------------------------------------------------------------------------------------------ AP bookkeeping
_SIZE_DELIMS = ((1 / 24) * (1 / 64), (1 / 4) * (1 / 12))  # utils/od_map.py:20-23


def _size_onehot(boxes, imsize):
    h, w = imsize
    area = (boxes[..., 2] - boxes[..., 0]) * (boxes[..., 3] - boxes[..., 1])
    lo, hi = _SIZE_DELIMS[0] * h * w, _SIZE_DELIMS[1] * h * w
    return torch.stack((torch.ones_like(area, dtype=torch.bool), area <= lo, (lo < area) & (area <= hi), hi < area), dim=2)


@torch.no_grad()
def prepare_od_map_stuffs(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk=50):
    """utils/od_map.py:214-287 as ONE kernel launch (fod_ap_match): top-50 per class, smoothed
    IoU, the 50-step greedy claim loop at 10 thresholds, size buckets.  Same outputs as the reference: confs
    (10,C+1,50B), is_positive, size_categories, num_annos.  Shapes outside the kernel's limits (more than 256
    detections or annotation slots) take the torch formulation below."""
    if pred_boxes.is_cuda and pred_boxes.shape[1] <= 256 and anno_boxes.shape[1] <= 256:
        return ops.ap_match(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk)
    return prepare_od_map_stuffs_torch(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk)


@torch.no_grad()
def prepare_od_map_stuffs_torch(pred_boxes, scores, anno_boxes, anno_classes, anno_active, imsize, topk=50):
    """The same bookkeeping with torch ops on the device (static shapes, no host sync): used beyond the kernel's size
    limits and as the kernel's on-device cross-check in tests/test_criterion_gpu.py."""
    dev = pred_boxes.device
    bsz, _, ncls = scores.shape
    thr = torch.arange(0.50, 1.00, 0.05, device=dev)
    # The reference first drops annotation slots that no sample uses (od_map.py:40-45, a data-dependent shape and a
    # host sync).  Inactive slots are masked out of `avail` below, so keeping all slots gives identical outputs with
    # static shapes — which lets the whole step be captured in a CUDA graph.
    ab, ac, aa = anno_boxes.float(), anno_classes, anno_active
    n = ab.shape[1]
    p, a = pred_boxes[:, :, None, :], ab[:, None, :, :]
    area_p = F.relu(p[..., 2] - p[..., 0]) * F.relu(p[..., 3] - p[..., 1])
    area_a = F.relu(a[..., 2] - a[..., 0]) * F.relu(a[..., 3] - a[..., 1])
    inter = F.relu(torch.min(p[..., 2], a[..., 2]) - torch.max(p[..., 0], a[..., 0])) * \
        F.relu(torch.min(p[..., 3], a[..., 3]) - torch.max(p[..., 1], a[..., 1]))
    iou = (inter + 1e-7) / (area_p + area_a - inter + 1e-7)
    conf, order = scores.sort(dim=1, descending=True)
    conf, order = conf[:, :topk], order[:, :topk]
    k = conf.shape[1]
    cls_ok = torch.cat((ac[:, None, :] == torch.arange(ncls - 1, device=dev)[None, :, None],
                        torch.ones((bsz, 1, n), dtype=torch.bool, device=dev)), dim=1)
    avail = (aa[:, None, :] == 1) & cls_ok
    iou_k = iou[:, :, None, :].expand(-1, -1, ncls, -1).gather(1, order[..., None].expand(-1, -1, -1, n))
    iou_k = torch.where(avail[:, None], iou_k, torch.zeros_like(iou_k))
    alive = torch.ones((bsz, len(thr), ncls, n), device=dev)
    positive = torch.zeros((bsz, len(thr), k, ncls), dtype=torch.bool, device=dev)
    for m in range(k):
        best, arg = (iou_k[:, None, m] * alive).max(dim=3)
        hit = best >= thr[None, :, None]
        positive[:, :, m] = hit
        cur = alive.gather(3, arg[..., None])
        alive.scatter_(3, arg[..., None], torch.where(hit[..., None], torch.zeros_like(cur), cur))
    confs = conf.reshape(bsz * k, ncls).t().reshape(1, ncls, bsz * k).repeat(len(thr), 1, 1)
    is_pos = positive.permute(1, 3, 0, 2).reshape(len(thr), ncls, bsz * k)
    sz = _size_onehot(pred_boxes, imsize)
    sz = sz[:, :, None, :].expand(-1, -1, ncls, -1).gather(1, order[..., None].expand(-1, -1, -1, 4))
    size_cat = sz.reshape(bsz * k, ncls, 4).permute(1, 2, 0)
    num_annos = (avail[..., None] & _size_onehot(ab, imsize)[:, None]).sum(dim=(0, 2))
    return confs, is_pos, size_cat, num_annos


# ------------------------------------------------------------------------------------------------ the wrapper
class SpatioTemporalDETR(nn.Module):
    """future_od/models/st_detr.py:58-234: same constructor, attributes (`_model`, `_criterion`), forward signature
    and return tuple `(outputs, state, loss, stats, od_map_stuffs)`."""

    def __init__(self, args, model, loss_matching_mode="per level"):
        super().__init__()
        self._model = model
        matcher = HungarianMatcher(args.set_cost_class, args.set_cost_bbox, args.set_cost_giou)
        weight_dict = {"loss_ce": args.cls_loss_coef, "loss_bbox": args.bbox_loss_coef, "loss_giou": args.giou_loss_coef}
        if args.aux_loss:
            for i in range(args.dec_layers - 1):
                weight_dict.update({f"{k}_{i}": v for k, v in list(weight_dict.items())[:3]})
        self._criterion = SetCriterion(args.num_classes, matcher, weight_dict, args.focal_alpha,
                                       ["labels", "boxes", "cardinality"], loss_matching_mode)
        self._imu_keys = ["translation", "acceleration", "rotation", "rotation_rate"]
        if not args.no_imu_speed:
            self._imu_keys.append("speed")
        self._encode_offset = args.encode_offset
        self.with_od_map = True
        self._whwh = {}

    @staticmethod
    def get_stat_idfs():
        return ["labels", "box_l1", "box_giou", "cardinality", "class_error"]

    def forward(self, data=None, visualize=False, epoch=None, distributed=False):
        images = data["video"]
        kwargs = {}
        if data.get("translation") is not None:
            kwargs["imu"] = torch.cat([data[k] for k in self._imu_keys], dim=2)
        if self._encode_offset:
            kwargs["temporal_offsets"] = data["temporal_offsets"]  # no numerical effect with no_temporal=True
        early = self._start_num_boxes(data) if (distributed and is_dist() and images.is_cuda and _EARLY_NUM_BOXES) else None
        outputs, moods = self._model(images, **kwargs)
        if not (isinstance(outputs, dict) and "pred_logits" in outputs and outputs["pred_logits"].dim() == 3):
            raise ValueError("cannot interpret output on the format: %s" % outputs)  # st_detr.py:156
        loss, stats = self.loss(data, outputs, distributed, early)
        od_map_stuffs, post = self.post_proc(outputs["pred_logits"][:, None], outputs["pred_boxes"][:, None], data, images)
        post["moods"] = moods
        return post, None, loss, stats, od_map_stuffs

    def _start_num_boxes(self, data):
        """The only data-path exchange of the loss (set_criterion.py:185-193: target count summed over ranks) depends on
        the batch alone, so it is issued BEFORE the forward pass, on a side stream: issued where the reference has it
        (inside the criterion) it sits on the main stream in the middle of the step and every rank waits there for the
        slowest one.  Returns the compacted targets and the reduced count; `loss` joins the side stream."""
        h, w = data["video"].shape[-2:]
        tb, tl, tc, tot = ops.targets_compact(data["boxes"], data["classes"], data["active"], h, w)
        n = tot.clone()
        if getattr(self, "_nb_stream", None) is None:
            self._nb_stream = torch.cuda.Stream()
        cur = torch.cuda.current_stream()
        self._nb_stream.wait_stream(cur)
        with torch.cuda.stream(self._nb_stream):
            torch.distributed.all_reduce(n)
        return tb, tl, tc, tot, n

    def loss(self, data, outputs, distributed, early=None):
        if early is not None:
            tb, tl, tc, tot, n = early
            torch.cuda.current_stream().wait_stream(self._nb_stream)
            num_boxes = (n / torch.distributed.get_world_size()).clamp(min=1.0)
            return self._criterion(outputs["all_logits"], outputs["all_boxes"], tb, tl, tc, tot, distributed, num_boxes)
        h, w = data["video"].shape[-2:]
        tb, tl, tc, tot = ops.targets_compact(data["boxes"], data["classes"], data["active"], h, w)
        return self._criterion(outputs["all_logits"], outputs["all_boxes"], tb, tl, tc, tot, distributed)

    @torch.no_grad()
    def post_proc(self, class_scores, boxes, data, images):
        """st_detr.py:190-234 for the one-output-frame case: class_scores (B,1,M,C), boxes (B,1,M,4)."""
        b, l, _, h, w = images.shape
        sc = torch.sigmoid(class_scores.float())
        sc = torch.cat((sc, sc.max(dim=3, keepdim=True)[0]), dim=3)
        key = (w, h, str(images.device))
        if key not in self._whwh:  # cached: a host->device constant copy is not allowed inside CUDA-graph capture
            self._whwh[key] = torch.tensor([w, h, w, h], device=images.device, dtype=torch.float32).view(1, 1, 1, 4)
        bx = boxes.float() * self._whwh[key]
        bx = torch.cat((bx[..., 0:2] - 0.5 * bx[..., 2:4], bx[..., 0:2] + 0.5 * bx[..., 2:4]), dim=-1)
        assert bx.shape[1] == 1, "If different #outs than #ins, #outs must be 1"
        stuffs = None
        if self.with_od_map:
            stuffs = prepare_od_map_stuffs(bx[:, 0], sc[:, 0], data["boxes"], data["classes"], data["active"], (h, w))
        return stuffs, {"class_scores": sc[:, :, None], "boxes": bx[:, :, None]}


# FOD_EARLY_NUM_BOXES=1 issues the num_boxes all-reduce before the forward pass on a side stream (_start_num_boxes).
# Measured at 8 GPUs on one box: device-timed step unchanged (28.44 / 28.46 vs 28.52 / 28.47 ms), end-to-end
# throughput WORSE (8 430 -> 7 070-7 740 clips/s: ranks that launch late now stall everyone at the first kernel), so
# the reference's placement inside the criterion stays the default.
_STEM_P4 = os.environ.get("FOD_STEM_P4", "1") != "0"  # 0: thread-assembled stem convolution (A/B switch)
_EARLY_NUM_BOXES = int(os.environ.get("FOD_EARLY_NUM_BOXES", "0")) != 0


def set_dropout(model, p):
    """Set the dropout probability of every transformer block (the reference hard-codes 0.1: runs/_model.py:42-55 and
    the layer defaults of transformer.py).  p = 0 gives the deterministic numerics the parity tests compare."""
    n = 0
    for m in model.modules():
        if isinstance(m, (M.EncoderAttention, M.EgodeepAttention, M.TransformerEncoderLayer, D.SlotToSlotAttention,
                          D.SlotToImageAttention, D.TransformerDecoderLayer)):
            m.p = float(p)
            n += 1
        elif isinstance(m, nn.Dropout):
            m.p = float(p)
    return n


def build_model(args, detr_args, dropout=None):
    """runs/_model.py:14-83 with the fod_b200 modules: same arguments, same module tree, same DDP wrapping.
    dropout=None keeps the reference's hard-coded 0.1 (runs/_model.py:52 and the encoder layer defaults);
    a number overrides every site (set_dropout) — 0.0 for deterministic parity runs."""
    lib.load()  # fail loudly before any weights are allocated if the CUDA library is missing
    d = detr_args.hidden_dim
    model = SpatioTemporalDETR(
        args=detr_args,
        model=FuturePredCore(
            separate_encoder=SeparateEncoder(
                backbone=M.CDetrBackbone(name="resnet50", train_backbone=detr_args.lr_backbone > 0, dilation=False,
                                         hidden_dim=d, pretrained=detr_args.pretrained_backbone),
                imu_layers=nn.Sequential(nn.Linear(14, 128), nn.ReLU(inplace=True), nn.Linear(128, d)),
                transformer=M.TransformerEncoder(layers=nn.ModuleList(
                    M.TransformerEncoderLayer(D=d, Nhead=detr_args.enc_nheads, Dff=detr_args.dim_feedforward,
                                              use_egodeep=True) for _ in range(detr_args.enc_layers))),
            ),
            joint_encoder=None,
            detector=D.CDetrDetectorSpatioTemporal(
                decoder=D.TransformerDecoder(
                    layers=nn.ModuleList([D.TransformerDecoderLayer(D=d, Nhead=detr_args.nheads,
                                                                    Dff=detr_args.dim_feedforward, dropout=0.1,
                                                                    num_images=2, use_slotstates=False)
                                          for _ in range(detr_args.dec_layers)]),
                    norm=nn.LayerNorm(d), return_intermediate=True, D=d),
                num_classes=detr_args.num_classes, hidden_dim=d, first_layer_special_when="always",
                num_queries=detr_args.num_queries, aux_loss=True, image_memory_mode="attend one at a time"),
            pos_encoder=M.PositionalEncoder(no_temporal=True),
        ),
    )
    if dropout is not None:
        set_dropout(model, dropout)
    model.to(args.device)
    if getattr(args, "distributed", False):
        model = nn.parallel.DistributedDataParallel(model, device_ids=[args.device], output_device=args.device,
                                                    find_unused_parameters=False)
    return model
