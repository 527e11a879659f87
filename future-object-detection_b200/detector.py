"""Decoder side of the drop-in model: ConditionalDETR-style decoder attending to the current and the previous frame.

Mirrors future_od/models/transformer.py:61-82 (SlotToSlotAttention), :122-181 (SlotToImageAttention), :184-312
(TransformerDecoderLayer), :315-398 (TransformerDecoder) and future_od/models/paper.py:280-429
(CDetrDetectorSpatioTemporal).  Queries are batch-first rows (B*M, 256) bf16.  `query_pos` is the same for every
sample (paper.py:370: query_embed.weight repeated), so every projection of it — and the reference points and their
sine embedding — is computed once on M rows and enters the batched GEMMs as a row-modular residual.
"""
import torch
import torch.nn as nn

from . import autograd as A
from . import ops
from .model import BF16, MLP, _lin, _lin_padded, _OutProjOnly, _q, _rows, _xavier_, gen_sineembed_for_position, \
    inverse_sigmoid


class SlotToSlotAttention(nn.Module):
    """transformer.py:61-82: q = Wqc·c + Wqp·p, k = Wkc·c + Wkp·p, v = Wv·c; MHA with head dim 32."""

    def __init__(self, D, Nhead, dropout=0.1):
        super().__init__()
        self.query_content = nn.Linear(D, D)
        self.query_pos = nn.Linear(D, D)
        self.key_content = nn.Linear(D, D)
        self.key_pos = nn.Linear(D, D)
        self.value = nn.Linear(D, D)
        self.fun = _OutProjOnly(D)
        self.heads = Nhead
        self.p = dropout

    def fused_param_groups(self):
        """Parameters read as one concatenated operand (optim.FlatTrainState gives them adjacent arena slots)."""
        return ([self.query_content.weight, self.key_content.weight, self.value.weight],
                [self.query_content.bias, self.key_content.bias, self.value.bias])

    def forward(self, c, qpos_bf, batch, m):
        d = c.shape[1]
        # batch-invariant positional halves (their biases ride along); content halves as ONE fused (3D x D) GEMM
        pq = _lin(qpos_bf, self.query_pos)
        pk = _lin(qpos_bf, self.key_pos)
        res = torch.cat((pq, pk, pq.new_zeros(m, d)), dim=1)
        gw, gb = self.fused_param_groups()
        self._fw = fw = A.FusedParams.build(gw, getattr(self, "_fw", None))
        self._fb = fb = A.FusedParams.build(gb, getattr(self, "_fb", None))
        if fw is not None and fb is not None:  # adjacent arena slots: no per-step cat, gradients written in place
            qkv = A.linear(c, fw, fb, fw._fod_shadow, residual=res, res_row_mod=m)
        else:
            w = torch.cat((self.query_content.weight, self.key_content.weight, self.value.weight), dim=0)
            b = torch.cat((self.query_content.bias, self.key_content.bias, self.value.bias), dim=0)
            wq = torch.cat((_q(self.query_content.weight), _q(self.key_content.weight), _q(self.value.weight)), dim=0)
            qkv = A.linear(c, w, b, wq, residual=res, res_row_mod=m)
        a = A.self_attention(qkv.view(batch, m, 3 * d), self.heads, (d // self.heads) ** -0.5,
                             drop=ops.drop_site(self.p, self.training))
        return a.view(-1, d)


class SlotToImageAttention(nn.Module):
    """transformer.py:122-181: per head q = [Wqc·c (+Wqp·p) ‖ Wqs·sine], k = [Wkc·x (+Wkp·pos) ‖ Wkp·pos], v = Wv·x,
    embed 2D (head dim 64), vdim D (head dim 32), scale 64^-0.5.  The concat is done inside the attention kernel."""

    def __init__(self, D, Nhead, dropout=0.1):
        super().__init__()
        self.query_content = nn.Linear(D, D)
        self.query_pos = nn.Linear(D, D)  # deleted by TransformerDecoder for layers > 0 (transformer.py:321-324)
        self.key_content = nn.Linear(D, D)
        self.key_pos = nn.Linear(D, D)
        self.value = nn.Linear(D, D)
        self.query_sine = nn.Linear(D, D)
        self.fun = _OutProjOnly(D)
        self.heads = Nhead
        self.p = dropout
        self.store_attention = False
        self.stored_attention = None

    def fused_param_groups(self):
        return ([self.key_content.weight, self.value.weight], [self.key_content.bias, self.value.bias])

    def forward(self, c, qpos_bf, qsine, img, img_pos_bf, is_first, batch, m, s):
        """c (B*M, D); qsine (B*M, D) or batch-invariant (M, D); img (B*S, D) tokens; img_pos_bf (S, D)."""
        d = c.shape[1]
        kp = _lin(img_pos_bf, self.key_pos)  # (S, D) batch-invariant key positional part
        gw, gb = self.fused_param_groups()
        self._fw = fw = A.FusedParams.build(gw, getattr(self, "_fw", None))
        self._fb = fb = A.FusedParams.build(gb, getattr(self, "_fb", None))
        if fw is not None and fb is not None:
            wkv, bkv, wkvq = fw, fb, fw._fod_shadow
        else:
            wkv = torch.cat((self.key_content.weight, self.value.weight), dim=0)
            bkv = torch.cat((self.key_content.bias, self.value.bias), dim=0)
            wkvq = torch.cat((_q(self.key_content.weight), _q(self.value.weight)), dim=0)
        if is_first:
            kv = A.linear(img, wkv, bkv, wkvq, residual=torch.cat((kp, kp.new_zeros(s, d)), dim=1), res_row_mod=s)
            qc = _lin(c, self.query_content, residual=_lin(qpos_bf, self.query_pos), res_row_mod=m)
        else:
            kv = A.linear(img, wkv, bkv, wkvq)
            qc = _lin(c, self.query_content)
        qs = _lin(qsine, self.query_sine)
        qs = qs.view(batch, m, d) if qs.shape[0] == batch * m else qs
        scale = (2 * d // self.heads) ** -0.5
        a = A.cross_attention(qc.view(batch, m, d), qs, kv.view(batch, s, 2 * d), kp, self.heads, scale,
                              drop=ops.drop_site(self.p, self.training))
        if self.store_attention:
            self.stored_attention = self._averaged_weights(qc.view(batch, m, d), qs, kv.view(batch, s, 2 * d), kp, scale)
        return a.view(-1, d)

    @torch.no_grad()
    def _averaged_weights(self, qc, qs, kv, kp, scale):
        """Head-averaged attention weights (B, M, S) for visualisation (transformer.py:179-180, demo.ipynb cell 9).
        Debug facility outside the hot path: plain torch on the projected operands."""
        b, m, d = qc.shape
        h, dh = self.heads, d // self.heads
        qs = qs.expand(b, m, d) if qs.dim() == 2 else qs
        q = torch.cat((qc.view(b, m, h, dh), qs.reshape(b, m, h, dh)), dim=3).float()
        k = torch.cat((kv[..., :d].reshape(b, -1, h, dh), kp.expand(b, -1, d).reshape(b, -1, h, dh)), dim=3).float()
        w = torch.einsum("bmhd,bshd->bhms", q * scale, k).softmax(dim=-1)
        return w.mean(dim=1)


class TransformerDecoderLayer(nn.Module):
    """transformer.py:184-312 with use_slotstates=False, use_egodeep=False (runs/_model.py:48-55)."""

    def __init__(self, D, Nhead, Dff, dropout=0.1, num_images=1, use_slotstates=False, use_egodeep=False):
        super().__init__()
        if use_slotstates or use_egodeep:
            raise NotImplementedError("slot-state / egodeep decoder variants are not built by runs/_model.py")
        self.self_attend = SlotToSlotAttention(D, Nhead, dropout)
        self.norm_sa = nn.LayerNorm(D)
        self.image_attend = nn.ModuleList(SlotToImageAttention(D, Nhead, dropout) for _ in range(num_images))
        self.norm_ia = nn.ModuleList(nn.LayerNorm(D) for _ in range(num_images))
        self.feedforward = nn.Sequential(nn.Linear(D, Dff), nn.ReLU(inplace=True), nn.Dropout(dropout),
                                         nn.Linear(Dff, D))
        self.norm_out = nn.LayerNorm(D)
        self.p = dropout

    def forward(self, c, qpos_bf, qsine, images, img_pos_bf, is_first, batch, m, s):
        site = lambda: ops.drop_site(self.p, self.training)  # noqa: E731 — dropout_sa / _ia / feedforward / _out
        sa = self.self_attend
        a = sa(c, qpos_bf, batch, m)
        op = sa.fun.out_proj
        c = A.linear_ln(a, c, op.weight, op.bias, _q(op.weight), self.norm_sa.weight, self.norm_sa.bias, self.norm_sa.eps,
                        drop=site())
        for i, img in enumerate(images):
            ia, nrm = self.image_attend[i], self.norm_ia[i]
            a = ia(c, qpos_bf, qsine, img, img_pos_bf, is_first, batch, m, s)
            op = ia.fun.out_proj
            c = A.linear_ln(a, c, op.weight, op.bias, _q(op.weight), nrm.weight, nrm.bias, nrm.eps, drop=site())
        l1, l2 = self.feedforward[0], self.feedforward[3]
        return A.ffn_ln(c, l1.weight, l1.bias, _q(l1.weight), l2.weight, l2.bias, _q(l2.weight), self.norm_out.weight,
                        self.norm_out.bias, self.norm_out.eps, drop_h=site(), drop_o=site())


class TransformerDecoder(nn.Module):
    """transformer.py:315-398."""

    def __init__(self, layers, norm=None, return_intermediate=False, D=256):
        super().__init__()
        self.layers = layers
        self.norm = norm
        self.return_intermediate = return_intermediate
        self.query_scale = MLP(D, D, D, 2)
        self.ref_point_head = MLP(D, D, 2, 2)
        for layer_id in range(1, len(layers)):  # transformer.py:321-324
            for ia in layers[layer_id].image_attend:
                ia.query_pos = None
        _xavier_(self)

    def forward(self, query_pos, images, img_pos_bf, batch, s, first_layer_special=True):
        """query_pos (M, D) fp32 (shared by the batch); images: list of (B*S, D) bf16 token matrices.

        Returns hs (layers, B*M, D) bf16 and the reference points (M, 2) fp32."""
        m, d = query_pos.shape
        qpos_bf = query_pos.to(BF16)
        ref = self.ref_point_head(qpos_bf).sigmoid()          # (M, 2) fp32, batch-invariant
        sine0 = gen_sineembed_for_position(ref, d)           # (M, D) fp32
        sine0_bf = sine0.to(BF16)
        c = torch.zeros((batch * m, d), dtype=BF16, device=query_pos.device)
        levels = []
        for layer_id, layer in enumerate(self.layers):
            special = layer_id == 0 and first_layer_special
            if special:
                qsine = sine0_bf
            else:
                scale = self.query_scale(c)
                qsine = (scale.view(batch, m, d) * sine0_bf).view(-1, d)
            c = layer(c, qpos_bf, qsine, images, img_pos_bf, special, batch, m, s)
            if self.return_intermediate:
                levels.append(A.add_ln(c, self.norm.weight, self.norm.bias, eps=self.norm.eps))
        if not self.return_intermediate:
            levels.append(A.add_ln(c, self.norm.weight, self.norm.bias, eps=self.norm.eps) if self.norm is not None else c)
        return torch.stack(levels), ref


class CDetrDetectorSpatioTemporal(nn.Module):
    """paper.py:280-429 for image_memory_mode="attend one at a time", use_slotstates=False."""

    def __init__(self, decoder, num_classes, hidden_dim, first_layer_special_when, num_queries, aux_loss=True,
                 image_memory_mode="attend one at a time", prune_dead_passes=True):
        super().__init__()
        if image_memory_mode != "attend one at a time":
            raise NotImplementedError(image_memory_mode)
        assert first_layer_special_when in ("always", "first frame", "never")
        self.decoder = decoder
        self.num_queries = num_queries
        self.aux_loss = aux_loss
        self.first_layer_special_when = first_layer_special_when
        self.num_images = len(decoder.layers[0].image_attend)
        self.class_embed = nn.Linear(hidden_dim, num_classes)
        self.bbox_embed = MLP(hidden_dim, hidden_dim, 4, 3)
        self.query_embed = nn.Embedding(num_queries, hidden_dim)
        import math
        nn.init.constant_(self.class_embed.bias, -math.log((1 - 0.01) / 0.01))  # paper.py:306-309
        nn.init.constant_(self.bbox_embed.layers[-1].weight, 0)                  # paper.py:312-313
        nn.init.constant_(self.bbox_embed.layers[-1].bias, 0)
        # Only the last frame's pass reaches the output (queries are re-zeroed every pass, paper.py:347-350,370-371;
        # SURVEY.md §0.3).  prune_dead_passes=False replays the T-1 dead passes as the reference does.
        self.prune_dead_passes = prune_dead_passes

    def forward(self, frames, img_pos_bf, batch, s):
        """frames: list over time of (B*S, D) bf16 token matrices (oldest first)."""
        out = None
        t = len(frames)
        for cur in (range(t) if not self.prune_dead_passes else [t - 1]):
            lst = [frames[cur - j] for j in range(self.num_images) if cur - j >= 0]
            special = self.first_layer_special_when == "always" or (self.first_layer_special_when == "first frame" and cur == 0)
            out = self.detect(lst, img_pos_bf, batch, s, special)
        return out

    def detect(self, images, img_pos_bf, batch, s, first_layer_special):
        """paper.py:352-419."""
        m = self.num_queries
        hs, ref = self.decoder(self.query_embed.weight, images, img_pos_bf, batch, s, first_layer_special)
        nl, _, d = hs.shape
        flat = hs.view(-1, d)
        ref_logit = inverse_sigmoid(ref)                                   # (M, 2)
        tmp = self.bbox_embed(flat).view(nl, batch, m, 4)                  # fp32 (padded head)
        tmp = torch.cat((tmp[..., :2] + ref_logit, tmp[..., 2:]), dim=-1)  # paper.py:410
        boxes = tmp.sigmoid()
        logits = _lin_padded(flat, self.class_embed).view(nl, batch, m, -1)
        out = {"pred_logits": logits[-1], "pred_boxes": boxes[-1], "hs": hs.view(nl, batch, m, d),
               "all_logits": logits, "all_boxes": boxes}
        if self.aux_loss:
            out["aux_outputs"] = [{"pred_logits": a, "pred_boxes": b} for a, b in zip(logits[:-1], boxes[:-1])]
        return out
