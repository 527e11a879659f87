"""Hammer for the attention forward on REAL operands: trains a few eager steps, grabs the first encoder attention call
(and the projection GEMM that feeds it), then replays [GEMM -> attention] chains inside CUDA graphs and counts launches
with a non-finite or deviating output.  Usage: python scripts/attn_hammer.py [replays] [chain]"""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
from fod_b200 import train as T
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict

replays = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
chain = int(sys.argv[2]) if len(sys.argv) > 2 else 16
WHICH = int(os.environ.get("WHICH", "0"))   # which attention call of the forward to grab
dev = "cuda:0"
grab = {"gemm": None, "attn": [], "on": False}
raw_gemm, raw_attn = ops.gemm, ops.attn_fwd


def g(*a, **kw):
    if grab["on"]:
        grab["gemm"] = (a, dict(kw))
    return raw_gemm(*a, **kw)


def f(*a, **kw):
    if grab["on"]:
        grab["attn"].append((a, dict(kw), grab["gemm"]))
    return raw_attn(*a, **kw)


ops.gemm, ops.attn_fwd = g, f
sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd); model.train()
opt = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay, wgrad_stream=False)
pool = [{k: v.to(dev) for k, v in make_clip(8, 3, 448, 800, seed=i).items()} for i in range(3)]
step = T.TrainStep(model, opt, args.max_norm)
for i in range(20):
    step(pool[i % 3])
grab["on"] = True
with torch.no_grad():
    model(data=pool[0], distributed=False)
grab["on"] = False
ops.gemm, ops.attn_fwd = raw_gemm, raw_attn
print("attention calls in one forward:", [(tuple(a[0].shape), a[0].stride(), tuple(a[1].shape)) for a, _, _ in grab["attn"]])
(aa, akw, gm) = grab["attn"][WHICH]
ga, gkw = gm
print("grabbed attention #%d: q1 %s stride %s, k1 %s, v %s, kw %s" % (WHICH, tuple(aa[0].shape), aa[0].stride(), tuple(aa[1].shape),
                                                                     tuple(aa[2].shape), {k: (tuple(v.shape) if torch.is_tensor(v) else v) for k, v in akw.items()}))
print("feeding gemm: x %s w %s kw %s" % (tuple(ga[0].shape), tuple(ga[1].shape), sorted(gkw)))
q1, k1, v = aa[0], aa[1], aa[2]
ref_o, ref_l = raw_attn(*aa, **akw)
print("reference finite:", bool(torch.isfinite(ref_o.float()).all()), bool(torch.isfinite(ref_l).all()),
      "lse range", float(ref_l.min()), float(ref_l.max()))
# does the feeding gemm write the tensor the attention reads?  (re-running it must reproduce q1's storage)
MODE = os.environ.get("MODE", "gemm")   # gemm: [gemm -> attention] chain; attn: attention only
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
outs = []
with torch.cuda.stream(side):
    for _ in range(2):
        if MODE == "gemm":
            raw_gemm(*ga, **gkw)
        raw_attn(*aa, **akw)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    for c in range(chain):
        if MODE == "gemm":
            keep = raw_gemm(*ga, **gkw)
        outs.append(raw_attn(*aa, **akw))
torch.cuda.synchronize()
bad = 0
for r in range(replays):
    graph.replay()
    flags = torch.stack([torch.stack([(o.float() - ref_o.float()).abs().max(), (l - ref_l).abs().max()]) for o, l in outs])
    fl = flags.cpu()
    if not bool(torch.isfinite(fl).all()) or float(fl.max()) > 0:
        bad += 1
        if bad <= 5:
            w = (~torch.isfinite(fl) | (fl > 0)).nonzero()
            print("replay %d: deviating launches %s  max dev %s" % (r, w.tolist()[:6], fl.max().item()))
            o, l = outs[w[0][0]]
            nf = (~torch.isfinite(l)).nonzero()
            dv = ((l - ref_l).abs() > 0).nonzero()
            print("   lse non-finite at (b,h,q):", nf.tolist()[:10], " count", nf.shape[0], " deviating count", dv.shape[0], dv.tolist()[:10])
print("mode %s: %d replays x %d launches, %d replays with a deviating launch" % (MODE, replays, chain, bad))
