import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict
dev = "cuda:0"
sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
data = {k: v.to(dev) for k, v in make_clip(2, 3, 128, 160, seed=3).items()}
model = build_model(Namespace(device=dev, distributed=False), dargs)
model.load_state_dict(sd); model.train()
flat = FlatTrainState(model, bucket_mb=8.0)
r = flat.reducer
count = {}
orig = r._hook
def hook(p):
    count[p] = count.get(p, 0) + 1
    orig(p)
r._hook = hook
for h in r._handles: h.remove()
r._handles = [p.register_post_accumulate_grad_hook(hook) for p in r.order]
flat.begin_step()
model(data)[2].backward()
flat.finish_backward()
names = {p: n for n, p in model.named_parameters()}
print("pending", r._pending)
for p in r.order:
    c = count.get(p, 0)
    if c != 1 or p._fod_uses != 0:
        print(names[p], "hook calls", c, "uses left", p._fod_uses)
