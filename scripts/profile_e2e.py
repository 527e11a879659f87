"""Where the end-to-end step loses time against the device-timed one: CUPTI trace of a few e2e iterations (pinned-host
batch -> prefetch, graph replay, ApAccumulator.add, loss read-back), reporting per iteration the graph span, the
kernels outside it and the idle gaps.  Usage: python scripts/profile_e2e.py [batch]"""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.od_map import ApAccumulator  # noqa: E402
from fod_b200.optim import FlatTrainState  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda:0")
torch.backends.cudnn.benchmark = True
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd)
model.train()
flat = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay)
host = [{k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in make_clip(B, 3, 448, 800, seed=i).items()} for i in range(3)]
devb = {k: v.to(dev) for k, v in host[0].items()}
eager = T.TrainStep(model, flat, args.max_norm)
for _ in range(3):
    eager(devb)
step = T.GraphedTrainStep(model, flat, devb, args.max_norm, warmup=1)
acc = ApAccumulator(distributed=False)
step.prefetch(host[0])


def it(i):
    loss = step()
    step.prefetch(host[(i + 1) % 3])
    acc.add(step.eager.od_map_stuffs, step.eager.stats)
    return float(loss.detach())


for i in range(5):
    it(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10):
    it(i)
e1.record()
torch.cuda.synchronize()
print("e2e ms/step: %.3f" % (e0.elapsed_time(e1) / 10))
e0.record()
for i in range(10):
    step(devb)
e1.record()
torch.cuda.synchronize()
print("device-resident ms/step: %.3f" % (e0.elapsed_time(e1) / 10))
step.prefetch(host[0]) if not step._ready else None
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    for i in range(4):
        it(i)
    torch.cuda.synchronize()
ev = []
for e in prof.profiler.kineto_results.events():
    if str(e.device_type()).endswith("CUDA") and e.duration_ns() > 0:
        ev.append((e.start_ns() / 1e3, e.duration_ns() / 1e3, e.device_resource_id(), e.name()))
ev.sort()
# iteration boundaries: the fused AdamW kernel ends a graph replay
ends = [s + d for s, d, st, n in ev if "adamw_flat" in n]
print("iterations seen:", len(ends))
for a, b in zip(ends[:-1], ends[1:]):
    seg = [(s, d, st, n) for s, d, st, n in ev if a <= s < b]
    h2d = [(s, d) for s, d, st, n in seg if "Memcpy HtoD" in n]
    d2d = [(s, d) for s, d, st, n in seg if "Memcpy DtoD" in n]
    first_graph = min(s for s, d, st, n in seg if "fod::" in n or "cutlass" in n or "cudnn" in n.lower())
    pre = [(s, d, n) for s, d, st, n in seg if s < first_graph and "Memcpy HtoD" not in n]
    busy_pre = sum(d for s, d, n in pre)
    print("period %.3f ms: tail+head before the graph's first kernel %.1f us (kernels/copies there: %d, busy %.1f us; DtoD %.1f us in %d copies); HtoD %.1f us total"
          % ((b - a) / 1e3, first_graph - a, len(pre), busy_pre, sum(d for s, d in d2d), len(d2d), sum(d for s, d in h2d)))
    for s, d, n in pre[:40]:
        print("     +%8.1f us  %7.1f us  %s" % (s - a, d, n[:70]))
    break
