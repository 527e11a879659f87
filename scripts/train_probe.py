"""Diagnostic: loss / activation-scale trajectory of the synthetic training run, eager vs CUDA graph.
Usage: python scripts/train_probe.py [graph] [steps]"""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

use_graph = "graph" in sys.argv
steps = int(sys.argv[-1]) if sys.argv[-1].isdigit() else 60
dev = "cuda:0"
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd)
model.train()
opt = T.setup_optimizer(model, args, capturable=use_graph)
pool = [{k: v.to(dev) for k, v in make_clip(8, 3, 448, 800, seed=i).items()} for i in range(3)]
step = T.TrainStep(model, opt, 0.1)
if "eagerfirst" in sys.argv:
    for i in range(8):
        step(pool[i % 3])
    torch.cuda.synchronize()
    print("eager-first done, status", int(model._criterion.last_status.item()))
if "pinned" in sys.argv:
    pool = [{k: v.cpu().pin_memory() for k, v in d.items()} for d in pool]
    dev0 = {k: v.to(dev) for k, v in pool[0].items()}
else:
    dev0 = pool[0]
if use_graph:
    step = T.GraphedTrainStep(model, opt, dev0, 0.1, warmup=1)
for i in range(steps):
    loss = step(pool[i % 3])
    st = int(model._criterion.last_status.item())
    if st or not bool(torch.isfinite(loss)):
        print("BAD at step", i, "status", st, "loss", float(loss))
        break
    if i % 50 == 0 or i == steps - 1:
        tok = model._model.last_encoder_tokens
        print(i, "loss", float(loss), "status", int(model._criterion.last_status.item()), "enc absmax", float(tok.float().abs().max()),
              "finite", bool(torch.isfinite(tok.float()).all()))
