"""BASELINE.json configs[2] and configs[4] on one GPU (configs[1] is bench.py's headline, configs[3] its --gpus N runs):
  * config 3  nusc_spatiotemporal_imu_250ms training step: 448x800 B=32 and the stage-2 shape 896x1600 B=16
  * config 5  nuim_spatiotemporal_imu inference sweep: T = 2..8 past frames, batch 1..64, per-frame encoder (drop-in
              mode, dead decoder passes pruned) and JointEncoder mode (one attention over all T*h*w tokens)
Every timing is a CUDA-graph replay bracketed by CUDA events after warm-up.  Prints one JSON object per line."""
import json
import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops  # noqa: E402
from fod_b200 import profile as prof  # noqa: E402
from fod_b200 import train as T  # noqa: E402
from fod_b200.optim import FlatTrainState  # noqa: E402
from fod_b200.st_detr import JointEncoder  # noqa: E402
from fod_b200.synthetic import make_clip, synthetic_state_dict  # noqa: E402

dev = "cuda:0"
sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)


def fresh(joint=False):
    m = build_model(Namespace(device=dev, distributed=False), dargs)
    m.load_state_dict(sd)
    if joint:
        se = m._model.separate_encoder
        m._model.joint_encoder = JointEncoder(se.transformer)
        se.transformer = None
    return m


def timed(fn, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def train_case(b, h, w, steps=10):
    model = fresh().train()
    flat = FlatTrainState(model, lr=dargs.lr, lr_backbone=dargs.lr_backbone, weight_decay=dargs.weight_decay)
    data = {k: v.to(dev) for k, v in make_clip(b, 3, h, w, seed=5).items()}
    step = T.GraphedTrainStep(model, flat, data, dargs.max_norm, warmup=2)
    for _ in range(3):
        step(data)
    ms = timed(lambda: step(data), steps)
    loss = float(step(data).detach())
    print(json.dumps({"config": "nusc_spatiotemporal_imu_250ms train step", "batch": b, "hw": [h, w], "ms_per_step": round(ms, 3),
                      "clips_per_s": round(b / ms * 1e3, 1), "loss": round(loss, 4), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2)}), flush=True)
    flat.close()
    del step, model, flat
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


def infer_case(model, joint, t, b, h=448, w=800, steps=10):
    data = {k: v.to(dev) for k, v in make_clip(b, t + 1, h, w, seed=6).items()}
    with torch.no_grad():
        for _ in range(2):
            model(data)
        prof.enable(True)
        model(data)
        kern = prof.disable()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            model(data)
    ms = timed(g.replay, steps)
    attn = [k for k in kern if k.startswith("attn_fwd_p1") and k.split("_")[-1].split("x")[0] not in ("128",)]
    iso = prof.isolated_us(attn)
    rec = {"config": "nuim_spatiotemporal_imu inference", "mode": "joint" if joint else "per-frame", "T": t, "batch": b,
           "ms_per_forward": round(ms, 3), "clips_per_s": round(b / ms * 1e3, 1)}
    for k in attn:
        if iso.get(k):
            r = kern[k]
            rec["encoder_" + k] = {"us": round(iso[k], 1), "tflops": round(r["flops"] / r["launches"] / (iso[k] * 1e-6) / 1e12, 1)}
    print(json.dumps(rec), flush=True)
    prof.clear_replays()
    del g
    torch.cuda.empty_cache()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "train"):
        train_case(32, 448, 800)
        train_case(16, 896, 1600)
    if which in ("all", "infer"):
        for joint in (False, True):
            model = fresh(joint).eval()
            for t in (2, 4, 8):
                for b in (1, 8, 64):
                    if joint and t * b > 128:
                        continue  # T*h*w = 2800 tokens x 64 clips: skipped to bound the sweep's run time
                    infer_case(model, joint, t, b)
            del model
            torch.cuda.empty_cache()
