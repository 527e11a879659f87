"""Op-level race tracer: repeats the SAME forward+backward (same clip, same parameters, no optimizer update) and
records a checksum of every fod_b200.ops call's outputs on the device (no host sync inside the step).  After each
repetition the checksums are compared with the first repetition's: the first op (in call order) whose checksum moves
beyond rounding is reported.  Ops built on floating-point atomics move in the last bits only; a race moves them grossly
or makes them non-finite.  Usage: python scripts/op_determinism.py [reps] [batch]   (env WS=1: weight-gradient stream)"""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
WS = int(os.environ.get("WS", "0"))
TOL = float(os.environ.get("TOL", "1e-3"))
dev = "cuda:0"
MAXOPS = 4096
sums = torch.zeros(MAXOPS, dtype=torch.float64, device=dev)
names = []
cursor = [0]
first_pass = [True]


def tensors(x):
    if isinstance(x, torch.Tensor):
        yield x
    elif isinstance(x, (tuple, list)):
        for y in x:
            yield from tensors(y)


def wrap(name, fn):
    def inner(*a, **kw):
        out = fn(*a, **kw)
        outs = list(tensors(out)) + [v for k, v in kw.items() if isinstance(v, torch.Tensor) and
                                     k in ("accum_into", "out", "colsum_out", "dzsum", "dgamma", "dbeta")]
        for t in outs:
            if not t.is_floating_point():
                t = t.double()
            i = cursor[0]
            cursor[0] += 1
            # weighted checksum (position-sensitive enough to see swapped or stale tiles)
            f = t.detach().reshape(-1).double()
            sums[i] = (f * f).sum() + f.sum()
            if first_pass[0]:
                names.append("%s#%d %s" % (name, i, tuple(t.shape)))
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
data = {k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=1).items()}
GRAPH = int(os.environ.get("GRAPH", "0"))
gsum_t = torch.zeros((), dtype=torch.float64, device=dev)


def one_rep():
    cursor[0] = 0
    opt.begin_step()
    _, _, loss, stats, _ = model(data=data, distributed=False)
    loss.backward()
    opt.finish_backward()
    g = opt.grads.double()
    gsum_t.copy_((g * g).sum())
    opt.grads.zero_()
    return loss


if GRAPH:
    # the same repetition as ONE CUDA graph (kernels back to back, fixed addresses): what GraphedTrainStep replays
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            one_rep()
            first_pass[0] = False
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        static_loss = one_rep()
    torch.cuda.synchronize()
    nops = cursor[0]

ref = None
ref_loss = ref_g = None
events = 0
inexact = {}
for r in range(reps):
    if GRAPH:
        sums.zero_()
        graph.replay()
        loss = static_loss
        cursor[0] = nops
    else:
        loss = one_rep()
    gsum = float(gsum_t)
    cur = sums[:cursor[0]].clone().cpu()
    lv = float(loss.detach())
    if ref is None:
        ref, ref_loss, ref_g = cur, lv, gsum
        first_pass[0] = False
        print("graph %d ws %d ops per step: %d, loss %.6f, |g|^2 %.6e" % (GRAPH, WS, cursor[0], lv, gsum), flush=True)
        continue
    rel = ((cur - ref).abs() / ref.abs().clamp_min(1e-30))
    rel[~torch.isfinite(cur)] = float("inf")
    bad = (rel > TOL).nonzero().flatten().tolist()
    for i in (rel > 0).nonzero().flatten().tolist():
        inexact[names[i].split("#")[0]] = max(inexact.get(names[i].split("#")[0], 0.0), float(rel[i]))
    if bad or lv != lv or abs(lv - ref_loss) > 1e-3 * abs(ref_loss) or abs(gsum - ref_g) > 1e-2 * ref_g or gsum != gsum:
        events += 1
        print("rep %d: loss %.6f (ref %.6f) |g|^2 %.6e (ref %.6e); ops off: %d" % (r, lv, ref_loss, gsum, ref_g, len(bad)))
        for i in bad[:6]:
            print("   first off: %s  checksum %.9e vs %.9e" % (names[i], float(cur[i]), float(ref[i])))
        sys.stdout.flush()
        if events >= 5:
            break
print("max relative checksum drift per op family (bit-inexact ops only):", {k: "%.2e" % v for k, v in inexact.items()})
print("repetitions %d, events %d" % (r + 1, events))
