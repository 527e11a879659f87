"""Locates the first non-finite op output inside graph-replayed TRAINING (optimizer included, clips cycling): every
fod_b200.ops call's outputs get a device-side checksum captured into the graph; after each replay the checksum vector
is read back and the first non-finite entry (call order) is reported together with the optimizer state.
Usage: python scripts/nan_graph.py [steps] [batch]   env WS=0/1"""
import json, os, sys
from argparse import Namespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import SpatioTemporalDETRArgs, build_model, ops
from fod_b200 import train as T
from fod_b200.optim import FlatTrainState
from fod_b200.synthetic import make_clip, synthetic_state_dict

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
WS = int(os.environ.get("WS", "1"))
dev = "cuda:0"
sums = torch.zeros(8192, dtype=torch.float64, device=dev)
names, cursor = [], [0]


def tensors(x):
    if isinstance(x, torch.Tensor):
        yield x
    elif isinstance(x, (tuple, list)):
        for y in x:
            yield from tensors(y)


SAVE = int(os.environ.get("SAVE", "0"))
saved = {}   # attention calls: persistent copies of the operands (the copies are captured into the graph)


def wrap(name, fn):
    def inner(*a, **kw):
        pending = None
        if name == "attn_fwd" and SAVE:
            key = cursor[0]
            ops_in = {"q1": a[0], "k1": a[1], "v": a[2], "q2": kw.get("q2"), "k2": kw.get("k2")}
            if key not in saved:
                saved[key] = {k: (torch.empty_like(t) if t is not None else None) for k, t in ops_in.items()}
                saved[key]["scale"], saved[key]["heads"], saved[key]["batch"] = a[3], kw.get("heads", 8), kw.get("batch")
            pending = (key, ops_in)
        ins = [t for t in list(tensors(a)) + [t for v in kw.values() for t in tensors(v)] if t.is_floating_point()]
        out = fn(*a, **kw)
        if pending is not None:   # operand copies AFTER the call: the kernel keeps its in-situ neighbours
            for k, t in pending[1].items():
                if t is not None:
                    saved[pending[0]][k].copy_(t)
            saved[pending[0]]["out"], saved[pending[0]]["lse"] = out[0].clone(), out[1].clone()
        outs = list(tensors(out)) + [v for k, v in kw.items() if isinstance(v, torch.Tensor) and
                                     k in ("accum_into", "out", "colsum_out", "dzsum", "dgamma", "dbeta")]
        i = cursor[0]
        cursor[0] += 2
        # entry i: inputs, entry i+1: outputs
        ins = [t for t in ins if t.is_cuda]
        outs = [t for t in outs if t.is_cuda]
        if ins:
            sums[i].copy_(torch.stack([t.detach().abs().max().double() for t in ins]).max())
        if outs:
            sums[i + 1].copy_(torch.stack([t.detach().abs().max().double() for t in outs]).max())
        if len(names) <= i:
            names.extend(["%s#%d inputs" % (name, i // 2), "%s#%d outputs %s" % (name, i // 2, [tuple(t.shape) for t in outs])])
        return out
    return inner


from fod_b200 import model as M
_raw_conv_bn = M._conv_bn
conv_idx = [0]


def _mon_conv_bn(x, conv, bn, relu, z=None):
    y = _raw_conv_bn(x, conv, bn, relu, z)
    i = cursor[0]
    cursor[0] += 2
    sums[i].copy_(x.detach().abs().max().double())
    sums[i + 1].copy_(y.detach().abs().max().double())
    if len(names) <= i:
        names.extend(["conv %s in%s k%s s%s z%d: input" % (tuple(conv.weight.shape), tuple(x.shape), conv.kernel_size, conv.stride, z is not None),
                      "conv %s z%d relu%d: output %s" % (tuple(conv.weight.shape), z is not None, relu, tuple(y.shape))])
    return y


M._conv_bn = _mon_conv_bn
TIGHT = float(os.environ.get("TIGHT", "1.3"))
RAW_ATTN = ops.attn_fwd
for name in ("gemm", "attn_fwd", "attn_bwd", "add_ln_fwd", "ln_bwd", "colsum", "maxpool3x3s2_nhwc", "targets_compact",
             "match", "set_loss"):
    setattr(ops, name, wrap(name, getattr(ops, name)))

sd = synthetic_state_dict(json.load(open(os.path.join(ROOT, "tests/golden/state_dict_template.json"))), seed=0, residual_gain=0.5)
args = SpatioTemporalDETRArgs(pretrained_backbone=False, num_classes=8, num_queries=128, lr_backbone=1e-4)
model = build_model(Namespace(device=dev, distributed=False), args)
model.load_state_dict(sd); model.train()
opt = FlatTrainState(model, lr=args.lr, lr_backbone=args.lr_backbone, weight_decay=args.weight_decay, wgrad_stream=bool(WS))
pool = [{k: v.to(dev) for k, v in make_clip(B, 3, 448, 800, seed=i).items()} for i in range(3)]


class Step(T.TrainStep):
    def __call__(self, data):
        cursor[0] = 0
        return super().__call__(data)


gs = T.GraphedTrainStep.__new__(T.GraphedTrainStep)
# same construction as GraphedTrainStep.__init__, with the cursor reset at the head of every (warm-up / captured) step
gs.eager = Step(model, opt, args.max_norm)
gs.static = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in pool[0].items()}
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        gs.eager(gs.static)
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
gs.graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(gs.graph):
    gs.loss = gs.eager(gs.static)
torch.cuda.synchronize()
nops = cursor[0]
print("graph captured: %d op checksums, ws %d" % (nops, WS), flush=True)
hist = []
for i in range(steps):
    loss = gs(pool[i % 3])
    cur = sums[:nops].cpu()
    st = opt.state.cpu().tolist()
    hist.append((i, float(loss), st[0], cur))
    hist = hist[-7:]
    if not bool(torch.isfinite(cur).all()) or not (st[0] < 2000.0):
        print("step %d: loss %s grad norm %s" % (i, float(loss), st[0]))
        print("history (step, loss, grad norm):", [(h[0], round(h[1], 3), h[2]) for h in hist])
        base = next((h for h in hist if h[0] == i - 3), hist[0])
        ratio = cur / base[3].clamp_min(1e-12)
        off = ((ratio > TIGHT) | ~torch.isfinite(cur)).nonzero().flatten().tolist()
        print("ops whose |x|max is > TIGHT x the value at step %d (call order), %d in total:" % (base[0], len(off)))
        for j in off[:30]:
            print("   [%d] %-60s %.4g  (was %.4g)" % (j, names[j][:60], float(cur[j]), float(base[3][j])))
        break
    if i % 100 == 0:
        print("step %d loss %.4f grad_norm %.4f" % (i, float(loss), st[0]), flush=True)
print("done")
