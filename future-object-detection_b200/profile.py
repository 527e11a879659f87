"""Per-kernel CUDA-event timing for bench.py's roofline report (measurement plumbing, off by default).

When enabled, every C-ABI call in ops.py is bracketed by a pair of CUDA events on the launching stream and tagged
with its ALGORITHMIC work (flops and bytes as defined in DESIGN.md §Measurement / SURVEY.md §8d).  `disable()`
synchronises and returns {kernel: {"ms", "launches", "flops", "bytes"}}.
"""
import collections

import torch

_on = False
_pending = []
_replay = {}  # kernel name -> (callable re-issuing the LAST launch of that name, objects it needs alive)


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


def end(start, name, flops=0.0, nbytes=0.0, replay=None, keep=None):
    if start is None:
        return
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    _pending.append((name, start, e, flops, nbytes))
    if replay is not None:
        _replay[name] = (replay, _storage_refs(keep))


def _storage_refs(obj):
    """Keep the MEMORY of the replayed launch alive without keeping its autograd graph alive (a stale AccumulateGrad
    node from an eager step would drag the default stream into a later CUDA-graph capture)."""
    if torch.is_tensor(obj):
        return obj.detach()
    if isinstance(obj, (tuple, list)):
        return [_storage_refs(o) for o in obj]
    return None


def isolated_us(names, per_graph=16, reps=8):
    """Device time per launch of the named kernels, each re-issued back to back inside a small CUDA graph (the launch
    the eager, CPU-launch-bound instrumented step brackets with events is dominated by launch gaps for kernels of a
    few microseconds; this is the warm, GPU-bound figure).  Returns {name: microseconds per launch}."""
    out = {}
    for name in names:
        if name not in _replay:
            continue
        fn, _keep = _replay[name]
        try:
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(2):
                    fn()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(per_graph):
                    fn()
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            out[name] = 1e3 * e0.elapsed_time(e1) / (reps * per_graph)
        except Exception as ex:  # measurement aid only: report instead of failing the bench
            out[name] = None
            print("isolated_us(%s) failed: %s" % (name, str(ex)[:200]))
    return out


def clear_replays():
    _replay.clear()


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
_BOUND = {"gemm": "tensor", "attn_fwd": "tensor", "attn_bwd": "tensor", "rows": "hbm", "loss": "hbm", "optim": "hbm"}


def _family(name):
    if name.startswith("gemm"):
        return "gemm"
    if name.startswith("attn_fwd"):
        return "attn_fwd"
    if name.startswith("attn_bwd"):
        return "attn_bwd"
    if name in ("match", "set_loss", "targets_compact", "ap_match"):
        return "loss"
    if name.startswith("adamw") or name.startswith("optim"):
        return "optim"
    return "rows"


def roofline_report(kern, peaks, total_ms_timed_region=None, traffic=None, isolated=None, steps=1, attention_s=None):
    """-> {"roofline": {...the (kernel, shape) with the largest time per step...}, "kernels": {per-kernel table}}.
    Fractions are of the MEASURED peaks (MEASURED_PEAKS.json) unless peaks["source"] == "fallback": the burst bf16
    figure for a launch timed alone (graph-replayed back to back), the sustained one for a launch timed inside the
    eager instrumented step.  attention_s: token count of the encoder's space-time attention; its forward and backward
    get their own entries under roofline["attention"] (the second half of BASELINE.json's metric)."""
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
    # Where an isolated (graph-replayed, GPU-bound) per-launch time is available it replaces the eager event bracket,
    # which for kernels of a few microseconds mostly measures the CPU launch gap.
    isolated = {k: v for k, v in (isolated or {}).items() if v and k in kern}
    for name, us in isolated.items():
        table[name]["us_per_launch_isolated"] = round(us, 2)
        r = kern[name]
        sec_iso = us * 1e-6 * r["launches"]
        table[name]["tflops_isolated"] = round(r["flops"] / sec_iso / 1e12, 2)
        table[name]["gbs_isolated"] = round(r["bytes"] / sec_iso / 1e9, 1)

    def per_step_ms(k):
        r = kern[k]
        return (isolated[k] * 1e-3 * r["launches"] if k in isolated else r["ms"]) / max(1, steps)

    for name in table:
        table[name]["ms_per_step"] = round(per_step_ms(name), 4)

    def entry(name):
        famname = _family(name)
        f = dict(kern[name])
        iso = name in isolated
        if iso:
            f["ms"] = isolated[name] * 1e-3 * f["launches"]
        sec = f["ms"] / 1e3
        tensor_peak = peaks["bf16_burst"] if iso else peaks["bf16_sustained"]
        # the bounding roofline of THIS shape: whichever of (algorithmic flops / tensor peak) and (algorithmic bytes /
        # HBM peak) takes longer; pure row / optimizer / loss kernels have no tensor work
        t_tensor = f["flops"] / (tensor_peak * 1e12)
        t_hbm = f["bytes"] / (peaks["hbm_gbs"] * 1e9)
        bound = "tensor" if (_BOUND[famname] == "tensor" and t_tensor >= t_hbm) else "hbm"
        if bound == "tensor":
            achieved, peak, unit = f["flops"] / sec / 1e12, tensor_peak, "TFLOP/s"
        else:
            achieved, peak, unit = f["bytes"] / sec / 1e9, peaks["hbm_gbs"], "GB/s"
        return {"kernel": name, "family": famname, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit,
                "frac": achieved / peak, "traffic": (traffic or {}).get(name), "peak_source": peaks["source"],
                "peak_kind": ("burst (launch timed alone)" if iso else "sustained (launch timed inside the step)"),
                "launches_per_step": f["launches"] / max(1, steps), "avg_launch_us": 1e3 * f["ms"] / max(1, f["launches"]),
                "ms_per_step": per_step_ms(name),
                "algorithmic_flops_per_launch": f["flops"] / max(1, f["launches"]),
                "algorithmic_bytes_per_launch": f["bytes"] / max(1, f["launches"]),
                "timing": "CUDA events around back-to-back graph replays of this launch" if iso
                else "CUDA events around each launch of the eager instrumented steps"}

    # The dominant kernel = the (kernel, shape) with the largest time per step among ALL instrumented launches.
    top_name = max(kern, key=per_step_ms)
    roof = entry(top_name)
    roof["ranked_by"] = "time per step over all %d instrumented (kernel, shape) pairs" % len(kern)
    roof["runners_up"] = [{"kernel": k, "ms_per_step": round(per_step_ms(k), 4)}
                          for k in sorted(kern, key=per_step_ms, reverse=True)[1:6]]
    if attention_s:
        att = {}
        for half in ("fwd", "bwd"):
            name = "attn_%s_p1_%dx%d" % (half, attention_s, attention_s)
            if name in kern:
                e = entry(name)
                # the tensor roofline is the metric's denominator ("ST-attention TFLOP/s vs bf16 peak") whatever bounds it
                sec = e["avg_launch_us"] * 1e-6
                e["tflops"] = e["algorithmic_flops_per_launch"] / sec / 1e12
                e["frac_of_bf16_peak"] = e["tflops"] / (peaks["bf16_burst"] if name in isolated else peaks["bf16_sustained"])
                att[half] = e
        roof["attention"] = att
    fams = {k: {"ms_total": round(v["ms"], 3), "launches": v["launches"],
                "tflops": round(v["flops"] / (v["ms"] / 1e3) / 1e12, 2) if v["ms"] > 0 else None,
                "gbs": round(v["bytes"] / (v["ms"] / 1e3) / 1e9, 1) if v["ms"] > 0 else None} for k, v in fam.items()}
    return {"roofline": roof, "kernel_families": fams, "kernels": table, "our_kernels_ms_in_instrumented_region": ours_ms}
