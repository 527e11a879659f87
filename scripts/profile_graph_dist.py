"""Timeline of ONE replay of the captured N-GPU training step (rank 0; torch.profiler / CUPTI): per-stream kernel
intervals, every NCCL kernel with its offset inside the step, and how long the optimizer waits for the gradient
all-reduce after the last backward kernel has finished (the exposed part of the collective).
Usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           scripts/profile_graph_dist.py [batch per GPU] [bucket MB] [dropout]"""
import collections
import json
import os
import sys
from argparse import Namespace

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.dist import shard_seed  # noqa: E402
from fod_b200.optim import FlatTrainState  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
BUCKET = float(sys.argv[2]) if len(sys.argv) > 2 else 32.0
P = float(sys.argv[3]) if len(sys.argv) > 3 else 0.1
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
torch.backends.cudnn.benchmark = True
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args, dropout=P)
model.load_state_dict(sd)
model.train()
flat = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay, bucket_mb=BUCKET)
data = {k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=shard_seed(rank, 0)).items()}
eager = T.TrainStep(model, flat, args.max_norm, distributed=world > 1)
for _ in range(3):
    eager(data)
step = T.GraphedTrainStep(model, flat, data, args.max_norm, warmup=1, distributed=world > 1)
for _ in range(3):
    step(data)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    step(data)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    step(data)
    torch.cuda.synchronize()
if world > 1:
    dist.barrier()
if rank == 0:
    ev = []
    for e in prof.profiler.kineto_results.events():
        if str(e.device_type()).endswith("CUDA") and e.duration_ns() > 0:
            ev.append((e.start_ns() / 1e3, e.duration_ns() / 1e3, e.device_resource_id(), e.name()))
    ev.sort()
    t0 = ev[0][0]
    t1 = max(s + d for s, d, _, _ in ev)
    print("N=%d  batch/GPU %d  bucket %g MB  dropout %g: graph step %.3f ms (events), profiled replay wall %.3f ms, %d device events"
          % (world, B, BUCKET, P, ms, (t1 - t0) / 1e3, len(ev)))
    per = collections.defaultdict(lambda: [0, 0.0, 1e30, 0.0])
    for s, d, st, n in ev:
        r = per[st]
        r[0] += 1
        r[1] += d
        r[2] = min(r[2], s - t0)
        r[3] = max(r[3], s + d - t0)
    print("per stream: id  kernels  busy ms  first start ms  last end ms")
    for st, r in sorted(per.items(), key=lambda kv: -kv[1][1]):
        print("  %4d %6d %9.3f %9.3f %9.3f" % (st, r[0], r[1] / 1e3, r[2] / 1e3, r[3] / 1e3))
    nccl = [(s - t0, d, st, n) for s, d, st, n in ev if "nccl" in n.lower()]
    opt = [(s - t0, d, n) for s, d, st, n in ev if "sumsq_partial" in n or "optim_prepare" in n or "adamw_flat" in n]
    print("NCCL kernels: %d, total %.3f ms" % (len(nccl), sum(d for _, d, _, _ in nccl) / 1e3))
    for s, d, st, n in nccl:
        print("  start %9.3f ms  dur %8.1f us  end %9.3f ms  stream %d  %s" % (s / 1e3, d, (s + d) / 1e3, st, n[:60]))
    if opt:
        o_start = opt[0][0]
        before = [(s - t0 + d, n) for s, d, st, n in ev if s - t0 < o_start and "nccl" not in n.lower()]
        last_compute_end, last_name = max(before)
        last_nccl_end = max([s + d for s, d, _, _ in nccl if s < o_start], default=0.0)
        print("optimizer starts at %.3f ms; last non-NCCL kernel before it ends at %.3f ms (%s); last NCCL kernel ends at %.3f ms"
              % (o_start / 1e3, last_compute_end / 1e3, last_name[:50], last_nccl_end / 1e3))
        print("exposed collective (optimizer start - end of last compute kernel): %.1f us" % (o_start - last_compute_end))
        print("optimizer kernels: " + ", ".join("%s %.1f us" % (n[:24], d) for _, d, n in opt))
sys.stdout.flush()
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
os._exit(0)
