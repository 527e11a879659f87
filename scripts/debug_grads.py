"""Diagnostic (not a test): per-parameter gradient comparison of the CUDA model against the fp32 CPU oracle on a
golden-sized clip.  Prints |grad|, |ref|, relative L2 error and cosine for every parameter group.
Usage on the GPU box: python scripts/debug_grads.py [fp32backbone]"""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402
from oracle import fod_oracle as O  # noqa: E402

dev = "cuda:0"
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
model = build_model(Namespace(device=dev, distributed=False),
                    SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4))
model.load_state_dict(sd)
model.train()
if "fp32backbone" in sys.argv:
    model._model.separate_encoder.backbone.compute_dtype = torch.float32
b, l, h, w = (2, 3, 128, 160) if "big" not in sys.argv else (2, 3, 448, 800)
data = make_clip(b, l, h, w, seed=1)
_, _, loss, _, _ = model({k: v.to(dev) for k, v in data.items()})
loss.backward()
ref_sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k else v) for k, v in sd.items()}
_, _, loss_ref, _, _, _ = O.forward(data, ref_sd, with_od_map=False)
loss_ref.backward()
print("loss", float(loss), float(loss_ref))
rows = []
for k, p in model.named_parameters():
    if p.grad is None or ref_sd[k].grad is None:
        continue
    g, r = p.grad.double().cpu().flatten(), ref_sd[k].grad.double().flatten()
    rel = float((g - r).norm() / max(1e-30, float(r.norm())))
    cos = float(torch.dot(g, r) / max(1e-30, float(g.norm() * r.norm())))
    rows.append((k, float(g.norm()), float(r.norm()), rel, cos))
for k, gn, rn, rel, cos in rows:
    flag = " <<<" if rel > 0.1 else ""
    print("%-95s |g| %.4e |ref| %.4e rel %.4f cos %.4f%s" % (k.replace("_model.", ""), gn, rn, rel, cos, flag))
k = "_model.separate_encoder.transformer.layers.0.self_attn.attn.in_proj_bias"
g, r = dict(model.named_parameters())[k].grad.cpu(), ref_sd[k].grad
for i, n in enumerate("qkv"):
    print("layer0 in_proj_bias", n, "mine", float(g[i * 256:(i + 1) * 256].norm()), "ref", float(r[i * 256:(i + 1) * 256].norm()),
          "err", float((g[i * 256:(i + 1) * 256] - r[i * 256:(i + 1) * 256]).norm()))
