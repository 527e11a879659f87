"""Full-size goldens at BASELINE.json's shapes, by the UNMODIFIED reference in the build container (CPU, fp32, eval):

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 \
      python tests/golden/make_golden_fullsize.py

  f1  configs[0]  nuim_spatiotemporal_imu forward: 1 clip x 4 frames (3 past), 448x800, S = 350 tokens per frame
  f2  configs[1]  shape (L = 3, 448x800), 2 clips: forward + loss + matcher indices + AP bookkeeping
  f3  configs[2]  stage-2 resolution (L = 3, 896x1600, S = 1400 tokens per frame), 1 clip: the size at which the
      space-time attention is tensor-bound           (python tests/golden/make_golden_fullsize.py f3_b1_l3_896x1600)

Weights: the name-keyed synthetic ones with residual_gain=0.5 — what bench.py trains.  Only the small
outputs are stored (the inputs are regenerated from the seed by fod_b200.synthetic.make_clip).
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

from future_od.models.st_detr import SpatioTemporalDETRArgs, to_detr_targets  # noqa: E402
from runs._model import build_model  # noqa: E402

CASES = {"f1_b1_l4_448x800": (1, 4, 448, 800, 21), "f2_b2_l3_448x800": (2, 3, 448, 800, 22),
         "f3_b1_l3_896x1600": (1, 3, 896, 1600, 23)}
IMU = ("translation", "acceleration", "rotation", "rotation_rate", "speed")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    model = build_model(Namespace(device="cpu", distributed=False),
                        SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4))
    with open(os.path.join(HERE, "state_dict_template.json")) as f:
        model.load_state_dict(synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5))
    model.eval()
    for name, (b, l, h, w, seed) in CASES.items():
        if len(sys.argv) > 1 and name not in sys.argv[1:]:
            continue
        data = make_clip(b, l, h, w, seed=seed)
        store = {}
        hook = model._model.detector.decoder.register_forward_hook(
            lambda m, i, o: store.__setitem__("hs", o[0].detach().clone()))
        with torch.no_grad():
            outputs, _, loss, stats, stuffs = model(data)
            core, _ = model._model(data["video"], imu=torch.cat([data[k] for k in IMU], dim=2))
        hook.remove()
        targets = to_detr_targets(H=h, W=w, anno_active=data["active"], anno_boxes=data["boxes"], anno_classes=data["classes"])
        levels = [core] + core["aux_outputs"]
        indices = [model._criterion.matcher({"pred_logits": lv["pred_logits"], "pred_boxes": lv["pred_boxes"]}, targets)
                   for lv in levels]
        torch.save({
            "case": (b, l, h, w, seed), "residual_gain": 0.5,
            "pred_logits": core["pred_logits"].clone(), "pred_boxes": core["pred_boxes"].clone(),
            "aux_logits": torch.stack([a["pred_logits"] for a in core["aux_outputs"]]),
            "aux_boxes": torch.stack([a["pred_boxes"] for a in core["aux_outputs"]]),
            "hs_last": store["hs"][-1].clone(), "loss": loss.detach().clone(),
            "stats": {k: v.detach().clone() for k, v in stats.items()},
            "class_scores": outputs["class_scores"].clone(), "boxes": outputs["boxes"].clone(),
            "od_map": [t.clone() for t in stuffs],
            "indices": [[(i.clone(), j.clone()) for i, j in lv] for lv in indices],
        }, os.path.join(HERE, name + ".pt"))
        print(name, "loss", float(loss), "size KB", os.path.getsize(os.path.join(HERE, name + ".pt")) // 1024)


if __name__ == "__main__":
    main()
