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
