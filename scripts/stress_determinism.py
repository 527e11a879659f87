"""Race detector: every kernel below is deterministic (no atomics on the tested paths), so repeated launches on the same
inputs must be bit-identical, also while a second stream keeps the SMs busy with other kernels.  Prints the number of
launches whose output differs from the first one.  Usage: python scripts/stress_determinism.py [iters]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import ops

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = "cuda:0"
torch.manual_seed(0)
BF = torch.bfloat16
side = torch.cuda.Stream()
noise_x = torch.randn(5600, 256, device=dev).to(BF)
noise_w = torch.randn(2048, 256, device=dev).to(BF)


def noise():
    with torch.cuda.stream(side):
        for _ in range(3):
            ops.gemm(noise_x, noise_w, relu=True)


def run(name, fn, concurrent):
    ref = [t.clone() for t in fn()]
    bad = nan = 0
    for i in range(iters):
        if concurrent:
            noise()
        outs = fn()
        same = all(torch.equal(a.view(torch.uint8), b.view(torch.uint8)) for a, b in zip(outs, ref))
        bad += (not same)
        nan += any(not bool(torch.isfinite(o.float()).all()) for o in outs)
    torch.cuda.synchronize()
    print("%-44s concurrent=%d  mismatching launches %d / %d  non-finite %d" % (name, concurrent, bad, iters, nan), flush=True)
    return bad


def attn_case(B, Lq, Lk, parts, bcast):
    C = 256
    q1 = torch.randn(B, Lq, C, device=dev).to(BF)
    k1 = torch.randn(B, Lk, C, device=dev).to(BF)
    v = torch.randn(B, Lk, C, device=dev).to(BF)
    q2 = k2 = None
    if parts == 2:
        q2 = torch.randn(B, Lq, C, device=dev).to(BF)
        k2 = torch.randn(Lk, C, device=dev).to(BF) if bcast else torch.randn(B, Lk, C, device=dev).to(BF)
    scale = (32 * parts) ** -0.5
    out, lse = ops.attn_fwd(q1, k1, v, scale, q2=q2, k2=k2)
    dout = torch.randn_like(out)
    fwd = lambda: ops.attn_fwd(q1, k1, v, scale, q2=q2, k2=k2)
    bwd = lambda: [t for t in ops.attn_bwd(q1, k1, v, out, lse, dout, scale, q2=q2, k2=k2) if t is not None]
    return fwd, bwd


total = 0
for (B, Lq, Lk, parts, bc) in [(24, 350, 350, 1, 0), (8, 1050, 1050, 1, 0), (8, 128, 350, 2, 1), (8, 128, 128, 1, 0),
                               (16, 1400, 1400, 1, 0), (24, 350, 350, 2, 1)]:
    fwd, bwd = attn_case(B, Lq, Lk, parts, bc)
    for conc in (0, 1):
        total += run("attn_fwd B%d %dx%d p%d bc%d" % (B, Lq, Lk, parts, bc), fwd, conc)
        total += run("attn_bwd B%d %dx%d p%d bc%d" % (B, Lq, Lk, parts, bc), bwd, conc)

x = torch.randn(8400, 256, device=dev).to(BF)
w1 = torch.randn(2048, 256, device=dev).to(BF) * 0.05
w2 = torch.randn(256, 2048, device=dev).to(BF) * 0.02
b1 = torch.randn(2048, device=dev)
b2 = torch.randn(256, device=dev)
gam, bet = torch.rand(256, device=dev) + 0.5, torch.randn(256, device=dev)
h = ops.gemm(x, w1, bias=b1, relu=True)
for conc in (0, 1):
    total += run("gemm relu 8400x2048x256", lambda: [ops.gemm(x, w1, bias=b1, relu=True)], conc)
    total += run("gemm ln 8400x256x2048", lambda: [ops.gemm(h, w2, bias=b2, residual=x, ln=(gam, bet, 1e-5))], conc)
    total += run("gemm 8400x768x256", lambda: [ops.gemm(x, noise_w[:768], bias=b1[:768])], conc)
print("TOTAL mismatches", total)
sys.exit(1 if total else 0)
