"""Golden vectors for the temporal sine code (future_od/models/paper.py:50-55,66-73), which
runs/_model.py:70-72 builds switched OFF (`PositionalEncoder(no_temporal=True)`).  Produced by the UNMODIFIED
reference modules in the build container with the switch flipped on the built model:

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 \
      python tests/golden/make_golden_temporal.py

Two cases: (t1) per-frame encoder, per-sample `temporal_offsets` (st_detr.py:116-117, `encode_offset=True`), full
forward + loss + matcher indices + gradient norms; (t2) no offsets (frame counter, paper.py:70) with the optional
JointEncoder (paper.py:180-203) sharing the per-frame encoder's layers, core outputs only.
"""
import json
import os
import sys
from argparse import Namespace

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import fod_b200  # noqa: E402,F401
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

from future_od.models.paper import JointEncoder  # noqa: E402
from future_od.models.st_detr import SpatioTemporalDETRArgs, to_detr_targets  # noqa: E402
from runs._model import build_model  # noqa: E402

IMU = ("translation", "acceleration", "rotation", "rotation_rate", "speed")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    model = build_model(Namespace(device="cpu", distributed=False),
                        SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128,
                                               lr_backbone=1e-4, encode_offset=True))
    with open(os.path.join(HERE, "state_dict_template.json")) as f:
        model.load_state_dict(synthetic_state_dict(json.load(f), seed=0))
    model.eval()
    pe = model._model.pos_encoder
    pe._no_temporal = False
    extra = float(pe._extra_temporal_offset)

    # ---- t1: offsets given, per-frame encoder, whole forward/backward
    b, l, h, w = 2, 3, 128, 160
    data = make_clip(b, l, h, w, seed=11)
    data["temporal_offsets"] = torch.tensor([[-1.0, -0.5, 0.0], [-0.75, -0.25, 0.0]])
    hs_store = {}
    hook = model._model.detector.decoder.register_forward_hook(
        lambda m, i, o: hs_store.__setitem__("hs", o[0].detach().clone()))
    model.zero_grad()
    outputs, _, loss, stats, stuffs = model(data)
    loss.backward()
    hook.remove()
    core_out, _ = model._model(data["video"], imu=torch.cat([data[k] for k in IMU], dim=2),
                               temporal_offsets=data["temporal_offsets"])
    targets = to_detr_targets(H=h, W=w, anno_active=data["active"], anno_boxes=data["boxes"], anno_classes=data["classes"])
    levels = [core_out] + core_out["aux_outputs"]
    indices = [model._criterion.matcher({"pred_logits": lv["pred_logits"], "pred_boxes": lv["pred_boxes"]}, targets)
               for lv in levels]
    grads = {k: p.grad.detach() for k, p in model.named_parameters() if p.grad is not None}
    torch.save({
        "case": (b, l, h, w, 11), "temporal_offsets": data["temporal_offsets"], "extra": extra,
        "pred_logits": core_out["pred_logits"].detach(), "pred_boxes": core_out["pred_boxes"].detach(),
        "aux_logits": torch.stack([a["pred_logits"] for a in core_out["aux_outputs"]]).detach(),
        "aux_boxes": torch.stack([a["pred_boxes"] for a in core_out["aux_outputs"]]).detach(),
        "hs_last": hs_store["hs"][-1].clone(), "loss": loss.detach(), "stats": {k: v.detach() for k, v in stats.items()},
        "indices": [[(i.clone(), j.clone()) for i, j in lv] for lv in indices],
        "grad_norms": {k: g.norm() for k, g in grads.items() if "backbone.body" not in k},
    }, os.path.join(HERE, "t1_temporal_b2_l3_128x160.pt"))
    print("t1 loss", float(loss), {k: float(v) for k, v in stats.items()})

    # ---- t2: frame-counter code + JointEncoder over all T*h*w tokens, core outputs only
    b, l, h, w = 2, 4, 96, 128
    data = make_clip(b, l, h, w, seed=12)
    model._model.joint_encoder = JointEncoder(model._model.separate_encoder.transformer)
    hook = model._model.detector.decoder.register_forward_hook(
        lambda m, i, o: hs_store.__setitem__("hs", o[0].detach().clone()))
    with torch.no_grad():
        core_out, _ = model._model(data["video"], imu=torch.cat([data[k] for k in IMU], dim=2))
    hook.remove()
    torch.save({
        "case": (b, l, h, w, 12), "extra": extra,
        "pred_logits": core_out["pred_logits"], "pred_boxes": core_out["pred_boxes"], "hs_last": hs_store["hs"][-1].clone(),
    }, os.path.join(HERE, "t2_temporal_joint_b2_l4_96x128.pt"))
    print("t2 logits mean", float(core_out["pred_logits"].mean()))


if __name__ == "__main__":
    main()
