"""The reference's OWN bf16 error: runs the unmodified reference on golden case g1 twice in the build container — fp32
and under torch.autocast(bfloat16) — and stores, per parameter tensor, the relative L2 distance between the two
gradients (plus the forward distances).  This is the yardstick for the bf16 gradient-parity gate of
tests/test_model_gpu.py: a bf16 implementation cannot be asked to sit closer to the fp32 reference than the
reference's own autocast path does.

  PYTHONPATH=/root/reference:/root/reference/ConditionalDETR PYTHONDONTWRITEBYTECODE=1 \
      python tests/golden/make_golden_autocast.py
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

from future_od.models.st_detr import SpatioTemporalDETRArgs  # noqa: E402
from runs._model import build_model  # noqa: E402

CASES = {"g1_b2_l3_128x160": (2, 3, 128, 160, 1, None), "g2_b1_l4_224x320": (1, 4, 224, 320, 2, None)}


def run(model, data, autocast):
    store = {}
    dec = model._model.detector.decoder
    hook = dec.register_forward_hook(lambda m, i, o: store.__setitem__("hs", o[0].detach().float().clone()))
    model.zero_grad()
    with torch.autocast("cpu", dtype=torch.bfloat16, enabled=autocast):
        outputs, _, loss, stats, _ = model(data)
    loss.backward()
    hook.remove()
    grads = {k: p.grad.detach().float().clone() for k, p in model.named_parameters() if p.grad is not None}
    return {"hs": store["hs"], "class_scores": outputs["class_scores"].detach().float(),
            "boxes": outputs["boxes"].detach().float(), "loss": float(loss), "grads": grads}


def rel(a, b):
    n = float(b.double().norm())
    return float((a.double() - b.double()).norm()) / n if n > 0 else float(a.double().norm())


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    model = build_model(Namespace(device="cpu", distributed=False),
                        SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4))
    model.load_state_dict(synthetic_state_dict(model.state_dict(), seed=0))
    model.eval()  # dropout off on both sides (the parity definition)
    out = {}
    for name, (b, l, h, w, seed, nt) in CASES.items():
        data = make_clip(b, l, h, w, seed=seed, fixed_targets=nt)
        f32 = run(model, data, False)
        b16 = run(model, data, True)
        g_all32 = torch.cat([g.flatten().double() for g in f32["grads"].values()])
        g_all16 = torch.cat([b16["grads"][k].flatten().double() for k in f32["grads"]])
        out[name] = {
            "forward": {k: rel(b16[k], f32[k]) for k in ("hs", "class_scores", "boxes")},
            "loss": [f32["loss"], b16["loss"]],
            "grad_global": float((g_all16 - g_all32).norm() / g_all32.norm()),
            "grad": {k: rel(b16["grads"][k], g) for k, g in f32["grads"].items()},
        }
        errs = sorted(out[name]["grad"].values())
        print(name, "forward", out[name]["forward"], "loss", out[name]["loss"], "grad global", out[name]["grad_global"],
              "per-tensor median %.4f p90 %.4f max %.4f" % (errs[len(errs) // 2], errs[int(0.9 * len(errs))], errs[-1]))
    with open(os.path.join(HERE, "ref_autocast_floor.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
