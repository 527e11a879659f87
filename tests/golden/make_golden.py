"""Generates tests/golden/*.pt by running the UNMODIFIED reference in the build container.

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 \
      python tests/golden/make_golden.py

The reference ships no golden vectors; these pin the oracle.  Weights are the
name-keyed synthetic ones (synthetic_state_dict), so only small outputs are stored.
"""
import json
import os
import sys
from argparse import Namespace

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import fod_b200  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

from future_od.models.st_detr import SpatioTemporalDETRArgs  # noqa: E402
from runs._model import build_model  # noqa: E402

CASES = {  # name: (B, L, H, W, clip seed, fixed_targets)
    "g1_b2_l3_128x160": (2, 3, 128, 160, 1, None),
    "g2_b1_l4_224x320": (1, 4, 224, 320, 2, None),
    "g3_b3_l3_96x128_nt0": (3, 3, 96, 128, 3, 0),
}


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    model = build_model(Namespace(device="cpu", distributed=False),
                        SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4))
    ref_sd = model.state_dict()
    with open(os.path.join(HERE, "state_dict_template.json"), "w") as f:
        json.dump({k: list(v.shape) for k, v in ref_sd.items()}, f)
    model.load_state_dict(synthetic_state_dict(ref_sd, seed=0))
    model.eval()
    for name, (b, l, h, w, seed, nt) in CASES.items():
        data = make_clip(b, l, h, w, seed=seed, fixed_targets=nt)
        hs_store = {}
        dec = model._model.detector.decoder
        hook = dec.register_forward_hook(lambda m, i, o: hs_store.__setitem__("hs", o[0].detach().clone()))
        enc_store = {}
        hook2 = model._model.separate_encoder.register_forward_hook(
            lambda m, i, o: enc_store.__setitem__("f", o[0].detach().clone()))
        model.zero_grad()
        outputs, _, loss, stats, stuffs = model(data)
        loss.backward()
        hook.remove()
        hook2.remove()
        core_out, _ = model._model(data["video"], imu=torch.cat(
            [data[k] for k in ("translation", "acceleration", "rotation", "rotation_rate", "speed")], dim=2))
        crit = model._criterion
        from future_od.models.st_detr import to_detr_targets
        targets = to_detr_targets(H=h, W=w, anno_active=data["active"], anno_boxes=data["boxes"], anno_classes=data["classes"])
        levels = [core_out] + core_out["aux_outputs"]
        indices = [crit.matcher({"pred_logits": lv["pred_logits"], "pred_boxes": lv["pred_boxes"]}, targets) for lv in levels]
        grads = {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}
        keep_full = [k for k in grads if grads[k].numel() <= 2048 and "backbone.body" not in k]
        gold = {
            "case": (b, l, h, w, seed, nt),
            "pred_logits": core_out["pred_logits"].detach(), "pred_boxes": core_out["pred_boxes"].detach(),
            "aux_logits": torch.stack([a["pred_logits"] for a in core_out["aux_outputs"]]).detach(),
            "aux_boxes": torch.stack([a["pred_boxes"] for a in core_out["aux_outputs"]]).detach(),
            "hs": hs_store["hs"], "encoder_features": enc_store["f"],
            "loss": loss.detach(), "stats": {k: v.detach() for k, v in stats.items()},
            "class_scores": outputs["class_scores"].detach(), "boxes": outputs["boxes"].detach(),
            "od_map": [t.detach() for t in stuffs],
            "indices": [[(i.clone(), j.clone()) for i, j in lv] for lv in indices],
            "grad_norms": {k: g.norm() for k, g in grads.items()},
            "grads_small": {k: grads[k] for k in keep_full[:60]},
        }
        torch.save(gold, os.path.join(HERE, name + ".pt"))
        print(name, "loss", float(loss), {k: float(v) for k, v in stats.items()},
              "size KB", os.path.getsize(os.path.join(HERE, name + ".pt")) // 1024)


if __name__ == "__main__":
    main()
