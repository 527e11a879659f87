"""Debug: train the small model eagerly and check every attention-backward result for non-finite values, comparing the
fused kernel with the two-kernel path on the same inputs when one shows up."""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
from fod_b200 import train as T
from fod_b200.synthetic import make_clip, synthetic_state_dict

dev = "cuda:0"
# pollute the allocator's free pool with NaN patterns, as earlier tests of a long pytest session do
junk = [torch.full((n,), float("nan"), device=dev) for n in (1 << 24, 1 << 22, 1 << 20, 1 << 18)]
junk += [torch.full((n,), float("nan"), device=dev, dtype=torch.bfloat16) for n in (1 << 24, 1 << 22, 1 << 20, 1 << 16, 1 << 14)]
del junk
orig = ops.attn_bwd
found = [0]

def checked(q1, k1, v, out, lse, dout, scale, heads=8, q2=None, k2=None, fused=None, mode=0):
    res = orig(q1, k1, v, out, lse, dout, scale, heads, q2=q2, k2=k2, fused=fused, mode=mode)
    bad = [i for i, t in enumerate(res) if t is not None and not bool(torch.isfinite(t.float()).all())]
    if bad and found[0] < 5:
        found[0] += 1
        ref = orig(q1, k1, v, out, lse, dout, scale, heads, q2=q2, k2=k2, fused=fused, mode=1)
        again = orig(q1, k1, v, out, lse, dout, scale, heads, q2=q2, k2=k2, fused=fused, mode=2)
        print("NON-FINITE attn_bwd outputs", bad, "shapes q", tuple(q1.shape), "k", tuple(k1.shape), "two", q2 is not None,
              "inputs finite", [bool(torch.isfinite(t.float()).all()) for t in (q1, k1, v, out, lse, dout)],
              "two-kernel finite", [bool(torch.isfinite(t.float()).all()) for t in ref if t is not None],
              "fused again finite", [bool(torch.isfinite(t.float()).all()) for t in again if t is not None])
        t = res[bad[0]].float()
        idx = (~torch.isfinite(t)).nonzero()
        print("  first bad indices", idx[:6].tolist(), "count", idx.shape[0], "of", t.numel())
    return res

ops.attn_bwd = checked
import fod_b200.autograd as A
A.ops.attn_bwd = checked
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0)
dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
pool = [{k: v.to(dev) for k, v in make_clip(2, 3, 128, 160, seed=30 + i).items()} for i in range(2)]
for rep in range(4):
    model = build_model(Namespace(device=dev, distributed=False), dargs)
    model.load_state_dict(sd)
    model.train()
    step = T.TrainStep(model, T.setup_optimizer(model, dargs), dargs.max_norm)
    losses = [float(step(pool[i % 2]).detach()) for i in range(24)]
    print("rep", rep, "losses", [round(x, 3) for x in losses[::6]], "finite", all(l == l for l in losses), "found", found[0])
