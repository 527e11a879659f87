"""First-non-finite tracer: runs eager training steps on the bench workload with every fod_b200.ops call checked for
non-finite outputs.  At the first hit it prints the op, whether its inputs were finite, re-runs it on the same inputs
(is it reproducible, i.e. numerics, or a race?) and stops.  Usage: python scripts/nan_trace.py [steps] [batch]"""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
from fod_b200 import train as T
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
WS = int(os.environ.get("WS", "0"))
dev = "cuda:0"
if int(os.environ.get("POISON", "0")):
    # torch.empty() returns NaN-filled memory: a kernel that reads what nobody wrote shows up deterministically
    torch.use_deterministic_algorithms(True, warn_only=True)
    torch.utils.deterministic.fill_uninitialized_memory = True
state = {"step": 0, "hit": False}


def tensors(x):
    if isinstance(x, torch.Tensor):
        yield x
    elif isinstance(x, (tuple, list)):
        for y in x:
            yield from tensors(y)
    elif isinstance(x, dict):
        for y in x.values():
            yield from tensors(y)


def finite(t):
    return (not t.is_floating_point()) or bool(torch.isfinite(t.float()).all())


def describe(t):
    f = t.float()
    ok = torch.isfinite(f)
    return "shape %s dtype %s finite %s nonfinite_count %d absmax(finite) %.4g" % (
        tuple(t.shape), str(t.dtype).replace("torch.", ""), bool(ok.all()), int((~ok).sum()),
        float(f[ok].abs().max()) if bool(ok.any()) else float("nan"))


def wrap(name, fn):
    def inner(*a, **kw):
        out = fn(*a, **kw)
        if state["hit"]:
            return out
        torch.cuda.synchronize()
        # in/out tensors: accum_into / out / colsum_out / dgamma ... are outputs too
        outs = list(tensors(out)) + [v for k, v in kw.items() if isinstance(v, torch.Tensor) and
                                     k in ("accum_into", "out", "colsum_out", "dzsum", "dgamma", "dbeta")]
        if all(finite(t) for t in outs):
            return out
        state["hit"] = True
        print("=== step %d: first non-finite output in ops.%s" % (state["step"], name), flush=True)
        for i, t in enumerate(tensors(a)):
            print("  arg tensor %d: %s" % (i, describe(t)))
        for k, v in kw.items():
            for t in tensors(v):
                print("  kw %s: %s" % (k, describe(t)))
            if not isinstance(v, torch.Tensor) and not isinstance(v, (tuple, list)):
                print("  kw %s = %r" % (k, v))
        for i, t in enumerate(outs):
            print("  out %d: %s" % (i, describe(t)))
        if not any(k in kw for k in ("accum_into", "colsum_out")):
            for r in range(3):
                again = fn(*a, **kw)
                torch.cuda.synchronize()
                print("  re-run %d: outputs finite %s" % (r, [finite(t) for t in tensors(again)]))
        return out
    return inner


for name in ("gemm", "attn_fwd", "attn_bwd", "add_ln_fwd", "ln_bwd", "colsum", "maxpool3x3s2_nhwc", "targets_compact",
             "match", "set_loss"):
    setattr(ops, name, wrap(name, getattr(ops, name)))

sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd); model.train()
opt = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay, wgrad_stream=bool(WS))
pool = [{k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=i).items()} for i in range(3)]
step = T.TrainStep(model, opt, args.max_norm)
for i in range(steps):
    state["step"] = i
    loss = step(pool[i % 3])
    pf = all(bool(torch.isfinite(p).all()) for p in model.parameters())
    if i % 10 == 0 or state["hit"] or not pf:
        print("step %3d loss %.4f grad_norm %.4f params_finite %s" % (i, float(loss.detach()), float(opt.state[0]), pf), flush=True)
    if state["hit"] or not pf:
        if not state["hit"]:
            print("parameters non-finite without any op output being non-finite: optimizer / torch-side op")
            gn = [n for n, p in model.named_parameters() if not bool(torch.isfinite(p).all())]
            print("  non-finite params:", gn[:20], len(gn))
        break
print("done, hit=%s" % state["hit"])
