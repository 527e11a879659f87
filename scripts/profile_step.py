"""Kernel-level time breakdown of one training step with torch.profiler (CUPTI) — an analysis aid; the committed
evidence is the ncu launch list under profiles/.  Usage: python scripts/profile_step.py [batch] [H] [W]"""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import setup_optimizer  # noqa: E402
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H = int(sys.argv[2]) if len(sys.argv) > 2 else 448
W = int(sys.argv[3]) if len(sys.argv) > 3 else 800
dev = "cuda:0"
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd)
model.train()
opt = setup_optimizer(model, args)
params = [p for p in model.parameters() if p.requires_grad]
data = {k: v.to(dev) for k, v in make_clip(B, 3, H, W, seed=0).items()}


def step():
    opt.zero_grad(set_to_none=True)
    _, _, loss, _, _ = model(data)
    loss.backward()
    torch.nn.utils.clip_grad_norm_(params, 0.1)
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(5):
    step()
torch.cuda.synchronize()
print("eager step ms:", 1e3 * (time.perf_counter() - t0) / 5)
# phase timing with events
ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
opt.zero_grad(set_to_none=True)
ev[0].record()
_, _, loss, _, _ = model(data)
ev[1].record()
loss.backward()
ev[2].record()
torch.nn.utils.clip_grad_norm_(params, 0.1)
ev[3].record()
opt.step()
ev[4].record()
torch.cuda.synchronize()
print("phases ms: fwd %.2f bwd %.2f clip %.2f adamw %.2f" % tuple(ev[i].elapsed_time(ev[i + 1]) for i in range(4)))
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=60, max_name_column_width=90))
