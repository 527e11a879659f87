"""Timeline of the CUDA-graph training step (torch.profiler / CUPTI): per-kernel warm durations, busy time vs wall time
and the idle gaps between consecutive kernels — tells whether the step is bound by kernel work or by node-to-node
latency.  Usage: python scripts/profile_graph.py [batch]"""
import collections
import json
import os
import re
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.optim import FlatTrainState  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = "cuda:0"
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd)
model.train()
flat = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay)
data = {k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=0).items()}
step = T.GraphedTrainStep(model, flat, data, args.max_norm, warmup=3)
for _ in range(3):
    step(data)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    step(data)
e1.record()
torch.cuda.synchronize()
print("graph step ms: %.3f" % (e0.elapsed_time(e1) / 10))
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    step(data)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
ev.sort(key=lambda e: e.time_range.start)
kern = [e for e in ev if "Memcpy" not in e.name and "Memset" not in e.name]
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
busy = sum(e.time_range.end - e.time_range.start for e in ev)
gaps, last_end = [], ev[0].time_range.end
for e in ev[1:]:
    if e.time_range.start > last_end:
        gaps.append(e.time_range.start - last_end)
    last_end = max(last_end, e.time_range.end)
print("events %d (kernels %d)  wall %.3f ms  busy(sum) %.3f ms  idle gaps %.3f ms over %d gaps (median %.2f us)" %
      (len(ev), len(kern), (t1 - t0) / 1e3, busy / 1e3, sum(gaps) / 1e3, len(gaps), sorted(gaps)[len(gaps) // 2] if gaps else 0))
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    n = re.sub(r"\(.*", "", e.name).replace("void ", "")[:90]
    agg[n][0] += 1
    agg[n][1] += e.time_range.end - e.time_range.start
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print("%-92s %5d %9.1f us %5.1f%% avg %6.2f" % (k, v[0], v[1], 100 * v[1] / busy, v[1] / v[0]))
