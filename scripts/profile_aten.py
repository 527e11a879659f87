"""Which torch (aten) ops are left in the step, with input shapes — an analysis aid for hunting element-wise passes.
Usage: python scripts/profile_aten.py [batch] [dropout]"""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.optim import FlatTrainState  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
P = float(sys.argv[2]) if len(sys.argv) > 2 else 0.1
dev = torch.device("cuda:0")
torch.backends.cudnn.benchmark = True
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args, dropout=P)
model.load_state_dict(sd)
model.train()
flat = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay)
step = T.TrainStep(model, flat, args.max_norm)
data = {k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=0).items()}
for _ in range(3):
    step(data)
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU],
                            record_shapes=True) as prof:
    step(data)
    torch.cuda.synchronize()
ka = prof.key_averages(group_by_input_shape=True)
rows = [e for e in ka if e.key.startswith("aten::") and e.self_device_time_total > 0 and "conv" not in e.key]
rows.sort(key=lambda e: -e.self_device_time_total)
print("aten (non-conv) self CUDA time: %.1f us over %d groups" % (sum(e.self_device_time_total for e in rows), len(rows)))
for e in rows[:45]:
    print("%8.1f us n=%3d %-28s %s" % (e.self_device_time_total, e.count, e.key, str(e.input_shapes)[:110]))
