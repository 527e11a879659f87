"""Per-kernel CUDA-event timing for bench.py's roofline report (measurement plumbing, off by default).

When enabled, every C-ABI call in ops.py is bracketed by a pair of CUDA events on the launching stream and tagged
with its ALGORITHMIC work (flops and bytes as defined in DESIGN.md §Measurement / SURVEY.md §8d).  `disable()`
synchronises and returns {kernel: {"ms", "launches", "flops", "bytes"}}.
"""
import collections

import torch

_on = False
_pending = []


def enable(flag=True):
    global _on
    _pending.clear()
    _on = bool(flag)


def begin():
    if not _on:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def end(start, name, flops=0.0, nbytes=0.0):
    if start is None:
        return
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    _pending.append((name, start, e, flops, nbytes))


def disable():
    global _on
    _on = False
    torch.cuda.synchronize()
    out = collections.defaultdict(lambda: {"ms": 0.0, "launches": 0, "flops": 0.0, "bytes": 0.0})
    for name, s, e, fl, nb in _pending:
        r = out[name]
        r["ms"] += s.elapsed_time(e)
        r["launches"] += 1
        r["flops"] += fl
        r["bytes"] += nb
    _pending.clear()
    return dict(out)


# which roofline bounds each kernel family
_BOUND = {"gemm": "tensor", "attn_fwd": "tensor", "attn_bwd": "tensor", "rows": "hbm", "loss": "hbm"}


def _family(name):
    if name.startswith("gemm"):
        return "gemm"
    if name.startswith("attn_fwd"):
        return "attn_fwd"
    if name.startswith("attn_bwd"):
        return "attn_bwd"
    if name in ("match", "set_loss", "targets_compact"):
        return "loss"
    return "rows"


def roofline_report(kern, peaks, total_ms_timed_region=None, traffic=None):
    """-> {"roofline": {...dominant kernel family...}, "kernels": {per-kernel table}}.  Fractions are of the MEASURED
    peaks (MEASURED_PEAKS.json) unless peaks["source"] == "fallback"; kernels timed inside a long step use the
    sustained bf16 figure."""
    if not kern:
        return {"roofline": None, "kernels": {}}
    fam = collections.defaultdict(lambda: {"ms": 0.0, "launches": 0, "flops": 0.0, "bytes": 0.0})
    table = {}
    for name, r in kern.items():
        f = fam[_family(name)]
        for k in ("ms", "launches", "flops", "bytes"):
            f[k] += r[k]
        sec = r["ms"] / 1e3
        table[name] = {"launches": r["launches"], "ms_total": round(r["ms"], 4),
                       "us_per_launch": round(1e3 * r["ms"] / max(1, r["launches"]), 2),
                       "tflops": round(r["flops"] / sec / 1e12, 2) if sec > 0 else None,
                       "gbs": round(r["bytes"] / sec / 1e9, 1) if sec > 0 else None}
    ours_ms = sum(f["ms"] for f in fam.values())
    top = max(fam, key=lambda k: fam[k]["ms"])
    f = fam[top]
    sec = f["ms"] / 1e3
    if _BOUND[top] == "tensor":
        achieved, peak, unit = f["flops"] / sec / 1e12, peaks["bf16_sustained"], "TFLOP/s"
    else:
        achieved, peak, unit = f["bytes"] / sec / 1e9, peaks["hbm_gbs"], "GB/s"
    roof = {"kernel": top, "bound": "tensor" if _BOUND[top] == "tensor" else "hbm", "achieved": achieved, "peak": peak,
            "unit": unit, "frac": achieved / peak, "traffic": (traffic or {}).get(top), "peak_source": peaks["source"],
            "launches": f["launches"], "avg_launch_us": 1e3 * f["ms"] / max(1, f["launches"]),
            "share_of_our_kernel_time": f["ms"] / ours_ms if ours_ms else None}
    fams = {k: {"ms_total": round(v["ms"], 3), "launches": v["launches"],
                "tflops": round(v["flops"] / (v["ms"] / 1e3) / 1e12, 2) if v["ms"] > 0 else None,
                "gbs": round(v["bytes"] / (v["ms"] / 1e3) / 1e9, 1) if v["ms"] > 0 else None} for k, v in fam.items()}
    return {"roofline": roof, "kernel_families": fams, "kernels": table, "our_kernels_ms_in_instrumented_region": ours_ms}
