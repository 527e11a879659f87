"""Debug: eager training of the small model; after every backward report the parameters whose gradient is non-finite."""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model
from fod_b200 import train as T
from fod_b200.synthetic import make_clip, synthetic_state_dict

dev = "cuda:0"
junk = [torch.full((n,), float("nan"), device=dev) for n in (1 << 24, 1 << 22, 1 << 20, 1 << 18)]
junk += [torch.full((n,), float("nan"), device=dev, dtype=torch.bfloat16) for n in (1 << 24, 1 << 22, 1 << 20, 1 << 16, 1 << 14)]
del junk
with open(os.path.join(ROOT, "tests", "golden", "state_dict_template.json")) as f:
    sd = synthetic_state_dict(json.load(f), seed=0, residual_gain=float(os.environ.get("GAIN", "1.0")))
dargs = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
pool = [{k: v.to(dev) for k, v in make_clip(2, 3, 128, 160, seed=30 + i).items()} for i in range(2)]
hits = 0
for rep in range(10):
    model = build_model(Namespace(device=dev, distributed=False), dargs)
    model.load_state_dict(sd)
    model.train()
    opt = T.setup_optimizer(model, dargs)
    for it in range(24):
        opt.zero_grad(set_to_none=True)
        _, _, loss, _, _ = model(data=pool[it % 2])
        loss.backward()
        bad = [n for n, p in model.named_parameters() if p.grad is not None and not bool(torch.isfinite(p.grad).all())]
        if bad:
            hits += 1
            total = sum(1 for p in model.parameters() if p.grad is not None)
            print("rep %d step %d loss %.4f: %d of %d gradients non-finite" % (rep, it, float(loss), len(bad), total))
            # group by module prefix
            import collections
            c = collections.Counter(".".join(n.split(".")[:6]) for n in bad)
            print("   ", c.most_common(12))
            good = [n for n, p in model.named_parameters() if p.grad is not None and bool(torch.isfinite(p.grad).all())]
            c2 = collections.Counter(".".join(n.split(".")[:6]) for n in good)
            print("    finite:", c2.most_common(12))
            break
        torch.nn.utils.clip_grad_norm_([p for p in model.parameters() if p.requires_grad], dargs.max_norm)
        opt.step()
    if hits >= 3:
        break
print("done, hits", hits)
