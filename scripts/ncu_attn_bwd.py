"""Driver for an ncu capture of the attention backward at the bench shape (B*T=16 frames, 350 tokens, 8 heads):
python scripts/ncu_attn_bwd.py  (run under ncu -k regex:attn_)"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fod_b200 import ops
B, L, C = 16, 350, 256
torch.manual_seed(0)
qkv = torch.randn(B, L, 3 * C, device="cuda").to(torch.bfloat16)
q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
scale = 32 ** -0.5
for _ in range(3):
    out, lse = ops.attn_fwd(q, k, v, scale)
    dout = torch.randn_like(out)
    ops.attn_bwd(q, k, v, out, lse, dout, scale, fused="qkv")
torch.cuda.synchronize()
if os.environ.get("TIME"):
    for (b, l) in [(16, 350), (16, 1400), (8, 1050)]:
        qkv = torch.randn(b, l, 3 * C, device="cuda").to(torch.bfloat16)
        q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
        out, lse = ops.attn_fwd(q, k, v, scale)
        dout = torch.randn_like(out)
        for _ in range(5):
            ops.attn_bwd(q, k, v, out, lse, dout, scale, fused="qkv")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                ops.attn_bwd(q, k, v, out, lse, dout, scale, fused="qkv")
        g.replay(); torch.cuda.synchronize()
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 20
        print("attn_bwd B%d S%d: %.1f us per call, %.1f TFLOP/s (10*B*S^2*256 algorithmic)" % (b, l, us, 10.0 * b * l * l * 256 / us / 1e6))
