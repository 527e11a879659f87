"""Run N graph-replayed training steps on the bench workload and print loss / gradient norm / matcher status every few
steps (divergence and NaN probe).  Usage: python scripts/train_probe.py [steps] [batch]"""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model
from fod_b200 import train as T
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = "cuda:0"
sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd); model.train()
FLAT, GRAPH, WS = (int(os.environ.get(k, "1")) for k in ("FLAT", "GRAPH", "WS"))
EVERY = int(os.environ.get("EVERY", "5"))
print("flat %d graph %d wgrad_stream %d" % (FLAT, GRAPH, WS))
if FLAT:
    opt = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay, wgrad_stream=bool(WS))
else:
    opt = T.setup_optimizer(model, args, capturable=bool(GRAPH))
pool = [{k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=i).items()} for i in range(3)]
step = T.GraphedTrainStep(model, opt, pool[0], args.max_norm, warmup=3) if GRAPH else T.TrainStep(model, opt, args.max_norm)
def attn_debug():
    import ctypes as C
    from fod_b200 import lib
    L = lib.load()
    if not hasattr(L, "fod_attn_debug_read"):
        return
    rec, n = (C.c_float * 128)(), C.c_int(0)
    L.fod_attn_debug_read(rec, C.byref(n))
    print("attention anomaly records: %d" % n.value)
    keys = "kind b h q j nkt m_used mx l o0 o1 s0 s1 extra block tid".split()
    for r in range(min(n.value, 8)):
        print("  ", {k: rec[r * 16 + i] for i, k in enumerate(keys)})


STOP = int(os.environ.get("STOP_ON_NAN", "0"))
hist = []
for i in range(steps):
    if STOP and FLAT:
        # state BEFORE the step (what the forward of this step reads)
        pa, sa = opt.params.abs(), opt.shadow.float().abs()
        hist.append("before step %d: |p|max %.4g at %d  |shadow|max %.4g at %d  |p-shadow|max %.4g  |m|max %.4g |v|max %.4g  prev norm %.4g clip %.4g" % (
            i, float(pa.max()), int(pa.argmax()), float(sa.max()), int(sa.argmax()),
            float((opt.params - opt.shadow.float()).abs().max()), float(opt.exp_avg.abs().max()),
            float(opt.exp_avg_sq.abs().max()), float(opt.state[0]), float(opt.state[1])))
    loss = step(pool[i % 3])
    if STOP and not bool(torch.isfinite(loss.detach())):
        print("step %d: non-finite loss" % i)
        print("\n".join(hist[-4:]))
        attn_debug()
        break
    if i % EVERY == 0 or i == steps - 1:
        torch.cuda.synchronize()
        finite = all(bool(torch.isfinite(p).all()) for p in model.parameters())
        extra = ("grad_norm %.4f clip %.4g" % (float(opt.state[0]), float(opt.state[1]))) if FLAT else ""
        print("step %3d loss %.4f %s match_status %d params_finite %s" % (
            i, float(loss.detach()), extra, int(model._criterion.last_status.item()), finite), flush=True)
