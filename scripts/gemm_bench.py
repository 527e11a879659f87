"""Per-launch times of the transformer GEMM shapes (B = 32 stage 1: 22400 token rows), each epilogue variant,
back-to-back launches in one CUDA graph (warm L2 where the operands fit).  usage: python scripts/gemm_bench.py"""
import json
import os
import sys

import torch

sys.path.insert(0, ".")
from fod_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
BF = torch.bfloat16


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (3 * reps)


M = int(sys.argv[1]) if len(sys.argv) > 1 else 22400
x256 = torch.randn(M, 256, device=dev).to(BF)
x2048 = torch.randn(M, 2048, device=dev).to(BF).relu()
w1 = (torch.randn(2048, 256, device=dev) * 0.05).to(BF)
w2 = (torch.randn(256, 2048, device=dev) * 0.05).to(BF)
wqkv = (torch.randn(768, 256, device=dev) * 0.05).to(BF)
wo = (torch.randn(256, 256, device=dev) * 0.05).to(BF)
b2048, b256, b768 = (torch.randn(n, device=dev) for n in (2048, 256, 768))
gam, bet = torch.ones(256, device=dev), torch.zeros(256, device=dev)
cs = torch.zeros(2048, device=dev)
ops.set_dropout_seed(1, dev)
cases = {
    "fwd 2048x256 bias+relu": lambda: ops.gemm(x256, w1, bias=b2048, relu=True),
    "fwd 2048x256 bias+relu+drop": lambda: ops.gemm(x256, w1, bias=b2048, relu=True, drop=(0.1, 3)),
    "fwd 2048x256 plain": lambda: ops.gemm(x256, w1),
    "dx 2048x256 mask+colsum": lambda: ops.gemm(x256, w2, b_mn=True, relu_mask=x2048, colsum_out=cs),
    "dx 2048x256 plain": lambda: ops.gemm(x256, w2, b_mn=True),
    "fwd 768x256 bias": lambda: ops.gemm(x256, wqkv, bias=b768),
    "fwd 256x256 bias": lambda: ops.gemm(x256, wo, bias=b256),
    "fwd_ln 256x2048": lambda: ops.gemm(x2048, w2, bias=b256, residual=x256, ln=(gam, bet, 1e-5)),
    "fwd_ln 256x2048 drop": lambda: ops.gemm(x2048, w2, bias=b256, residual=x256, ln=(gam, bet, 1e-5), drop=(0.1, 4)),
    "fwd_ln 256x256": lambda: ops.gemm(x256, wo, bias=b256, residual=x256, ln=(gam, bet, 1e-5)),
    "dx 256x2048 residual": lambda: ops.gemm(x2048, w1, b_mn=True, residual=x256),
    "dx 256x2048 plain": lambda: ops.gemm(x2048, w1, b_mn=True),
}
# attention / element-wise cases (one encoder layer at B = 32: 64 sequences of 350 tokens; stage 2: S = 1400)
def _attn_case(nseq, S, drop, bwd, mode=0, bits=False):
    st = {}
    d = (0.1, 9) if drop else None
    if bits:  # a {seed, step} snapshot in the tuple: keep decisions drawn once (fod_attn_keep_bits) and cached on it
        d = (0.1, 9, ops.rng_state(dev).clone())

    def run():
        if not st:  # built on first use, so a GEMM_CASE filter launches nothing else
            qkv = torch.randn(nseq, S, 768, device=dev).to(BF)
            st["q"], st["k"], st["v"] = qkv[..., :256], qkv[..., 256:512], qkv[..., 512:]
            if bwd:
                st["o"], st["lse"] = ops.attn_fwd(st["q"], st["k"], st["v"], 32 ** -0.5, 8, drop=d)
                st["do"] = torch.randn_like(st["o"])
        if bwd:
            return ops.attn_bwd(st["q"], st["k"], st["v"], st["o"], st["lse"], st["do"], 32 ** -0.5, 8, mode=mode, drop=d)
        return ops.attn_fwd(st["q"], st["k"], st["v"], 32 ** -0.5, 8, drop=d)
    return run


def _relu_case(n):
    st = {}

    def run():
        if not st:
            st["t"] = [torch.randn(n, device=dev).to(BF) for _ in range(3)]
        return ops.relu_bwd_add(*st["t"])
    return run


cases["fwd_ln_save 256x2048 drop"] = lambda: ops.gemm(x2048, w2, bias=b256, residual=x256, ln=(gam, bet, 1e-5), save_ln=True, drop=(0.1, 4))
# weight gradients: dW (N, K) += dYᵀ (N, rows) · X (rows, K), both operands MN-major, split-K with fp32 reductions
_gw1, _gw2 = torch.zeros(2048, 256, device=dev), torch.zeros(256, 2048, device=dev)
_gqkv, _go = torch.zeros(768, 256, device=dev), torch.zeros(256, 256, device=dev)
_x768 = torch.randn(M, 768, device=dev).to(BF)
cases["dw 2048x256"] = lambda: ops.gemm(x2048, x256, a_mn=True, b_mn=True, accum_into=_gw1)
cases["dw 256x2048"] = lambda: ops.gemm(x256, x2048, a_mn=True, b_mn=True, accum_into=_gw2)
cases["dw 768x256"] = lambda: ops.gemm(_x768, x256, a_mn=True, b_mn=True, accum_into=_gqkv)
cases["dw 256x256"] = lambda: ops.gemm(x256, x256, a_mn=True, b_mn=True, accum_into=_go)
def _ln_bwd_case(rows, drop):
    st = {}

    def run():
        if not st:
            st["dy"], st["pre"] = (torch.randn(rows, 256, device=dev).to(BF) for _ in range(2))
            st["mean"], st["rstd"] = torch.zeros(rows, device=dev), torch.ones(rows, device=dev)
            st["acc"] = torch.zeros(3, 256, device=dev)
        a = st["acc"]
        return ops.ln_bwd(st["dy"], st["pre"], st["mean"], st["rstd"], gam, a[0], a[1], dzsum=a[2], drop=(0.1, 5) if drop else None)
    return run


cases["ln_bwd 22400"] = _ln_bwd_case(22400, False)
cases["ln_bwd 22400 drop"] = _ln_bwd_case(22400, True)
cases["ln_bwd 4096 drop"] = _ln_bwd_case(4096, True)
cases["ln_bwd 128 drop"] = _ln_bwd_case(128, True)
cases["attn_fwd 64x350 drop"] = _attn_case(64, 350, True, False)
cases["attn_fwd 64x350"] = _attn_case(64, 350, False, False)
cases["attn_bwd 64x350 drop"] = _attn_case(64, 350, True, True)
cases["attn_bwd 64x350"] = _attn_case(64, 350, False, True)
cases["attn_fwd 64x350 drop bits"] = _attn_case(64, 350, True, False, bits=True)
cases["attn_bwd 64x350 drop bits"] = _attn_case(64, 350, True, True, bits=True)


def _bits_case(rows, lk):
    st = {}

    def run():
        if not st:
            st["o"] = torch.empty((rows, 16 * ((lk + 127) // 128)), dtype=torch.uint8, device=dev)
            st["s"] = ops.rng_state(dev).clone()
        from fod_b200 import lib as _lib
        import ctypes as _C
        L = _lib.load()
        return L.fod_attn_keep_bits(_C.c_void_p(st["s"].data_ptr()), 9, 0.1, rows, lk, _C.c_void_p(st["o"].data_ptr()), st["o"].shape[1],
                                    _C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return run


cases["keep_bits 64x8x350x350"] = _bits_case(64 * 8 * 350, 350)
cases["attn_fwd 32x1400"] = _attn_case(32, 1400, False, False)
cases["attn_bwd 32x1400"] = _attn_case(32, 1400, False, True)
cases["relu_bwd_add3 183500800"] = _relu_case(183500800)


def _stem_case(p4):
    st = {}

    def run():
        if not st:
            st["v"] = torch.randn(64, 1, 3, 448, 800, device=dev)
            w, s_ = torch.randn(64, 3, 7, 7, device=dev) * 0.1, torch.rand(64, device=dev) + 0.5
            st["t"] = torch.randn(64, device=dev)
            if p4:
                st["x"], st["w"] = ops.frames_to_nhwc4p_bf16(st["v"]), ops.stem_weights4(w, s_)
            else:
                st["x"], st["w"] = ops.frames_to_nhwc_bf16(st["v"]), ops.stem_weights(w, s_)
        return (ops.stem_conv7x7s2_p4 if p4 else ops.stem_conv7x7s2)(st["x"], st["w"], st["t"])
    return run


def _frames_case(p4):
    st = {}

    def run():
        if not st:
            st["v"] = torch.randn(32, 3, 3, 448, 800, device=dev)
        return (ops.frames_to_nhwc4p_bf16 if p4 else ops.frames_to_nhwc_bf16)(st["v"], frames=2)
    return run


cases["stem 64x448x800 threads"] = _stem_case(False)
cases["stem 64x448x800 tma"] = _stem_case(True)


def _stem_pool_case(fused):
    st = {}

    def run():
        if not st:
            st["v"] = torch.randn(64, 1, 3, 448, 800, device=dev)
            w, s_ = torch.randn(64, 3, 7, 7, device=dev) * 0.1, torch.rand(64, device=dev) + 0.5
            st["t"] = torch.randn(64, device=dev)
            st["x"], st["w"] = ops.frames_to_nhwc4p_bf16(st["v"]), ops.stem_weights4(w, s_)
        if fused:
            return ops.stem_conv_pool_p4(st["x"], st["w"], st["t"])
        return ops.maxpool3x3s2_nhwc(ops.stem_conv7x7s2_p4(st["x"], st["w"], st["t"]))
    return run


cases["stem+pool 64x448x800 two kernels"] = _stem_pool_case(False)
cases["stem+pool 64x448x800 fused"] = _stem_pool_case(True)
cases["frames 64x448x800 nhwc3"] = _frames_case(False)
cases["frames 64x448x800 nhwc4p"] = _frames_case(True)

flops = {"2048x256": 2.0 * M * 2048 * 256, "768x256": 2.0 * M * 768 * 256, "256x256": 2.0 * M * 256 * 256,
         "256x2048": 2.0 * M * 2048 * 256}
only = os.environ.get("GEMM_CASE")
match = os.environ.get("GEMM_MATCH")
for name, fn in cases.items():
    if only and only != name:
        continue
    if match and not any(m in name for m in match.split(",")):  # GEMM_MATCH: comma-separated substrings
        continue
    if os.environ.get("GEMM_EAGER"):  # under ncu: plain launches, no graph
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        continue
    us = timed(fn)
    f = flops.get(name.split()[1])
    if name.startswith("dw"):
        nn, kk = (int(v) for v in name.split()[1].split("x"))
        f = 2.0 * M * nn * kk
    if f is None and name.startswith("attn"):
        nseq, S = (int(v) for v in name.split()[1].split("x"))
        f = (4.0 if "fwd" in name else 10.0) * nseq * S * S * 256
    print(json.dumps({"case": name, "M": M, "us": round(us, 2), "tflops": round(f / us / 1e6, 1) if f else None}))
