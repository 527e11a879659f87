"""Stock comparator for piece (1): torch.nn.functional.scaled_dot_product_attention (library flash /
cuDNN kernels, bf16) against fod_attn_fwd / fod_attn_bwd on the encoder's attention shapes.  Head dim 32, 8 heads.
Run on a B200:  python scripts/sdpa_compare.py > gpurun_out/sdpa_compare.jsonl
Timing: CUDA events around back-to-back launches after warm-up; flops = 4·N·S²·D forward, 10·N·S²·D backward."""
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fod_b200 import ops  # noqa: E402

DEV, H, DH = "cuda:0", 8, 32
D = H * DH


def timed(fn, n=50, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us


def main():
    g = torch.Generator(device=DEV).manual_seed(0)
    for n_seq, s in ((16, 350), (64, 350), (16, 1400), (32, 1400), (8, 2800), (8, 5600)):
        qkv = torch.randn(n_seq, s, 3 * D, device=DEV, dtype=torch.bfloat16, generator=g)
        q, k, v = qkv[..., :D], qkv[..., D:2 * D], qkv[..., 2 * D:]
        scale = DH ** -0.5
        out, lse = ops.attn_fwd(q, k, v, scale, H)
        dout = torch.randn_like(out)
        # (N, S, H, dh) -> (N, H, S, dh) views for the library call
        qh, kh, vh = (t.reshape(n_seq, s, H, DH).transpose(1, 2) for t in (q, k, v))
        ref = F.scaled_dot_product_attention(qh, kh, vh, scale=scale).transpose(1, 2).reshape(n_seq, s, D)
        err = float((out.float() - ref.float()).norm() / ref.float().norm())
        fwd_flops, bwd_flops = 4.0 * n_seq * s * s * D, 10.0 * n_seq * s * s * D
        row = {"n_seq": n_seq, "S": s, "rel_l2_vs_sdpa": round(err, 5)}
        t = timed(lambda: ops.attn_fwd(q, k, v, scale, H))
        row["fod_fwd_us"], row["fod_fwd_tflops"] = round(t, 1), round(fwd_flops / t * 1e-6, 1)
        t = timed(lambda: F.scaled_dot_product_attention(qh, kh, vh, scale=scale))
        row["sdpa_fwd_us"], row["sdpa_fwd_tflops"] = round(t, 1), round(fwd_flops / t * 1e-6, 1)
        t = timed(lambda: ops.attn_bwd(q, k, v, out, lse, dout, scale, H, fused="qkv"))
        row["fod_bwd_us"], row["fod_bwd_tflops"] = round(t, 1), round(bwd_flops / t * 1e-6, 1)
        ql, kl, vl = (t_.detach().clone().requires_grad_(True) for t_ in (qh, kh, vh))
        o2 = F.scaled_dot_product_attention(ql, kl, vl, scale=scale)
        do2 = dout.reshape(n_seq, s, H, DH).transpose(1, 2)
        t = timed(lambda: torch.autograd.grad(o2, (ql, kl, vl), do2, retain_graph=True))
        row["sdpa_bwd_us"], row["sdpa_bwd_tflops"] = round(t, 1), round(bwd_flops / t * 1e-6, 1)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
