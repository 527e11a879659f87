#!/bin/bash
# usage: r2_bench.sh [extra bench args]  — default bench line + top kernels
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 "$@" > $O/bench.json 2> $O/bench.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "gpu_launches", "clocks")})
    r = dict(d["roofline"]); att = r.pop("attention", None); print(json.dumps(r)[:1200]); print("attention:", json.dumps(att)[:1500])
    ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"])[:40]
    for k, v in ks: print("%-34s n=%4d ms/step %.4f iso_us %s tf %s gbs %s" % (k, v["launches"], v["ms_per_step"], v.get("us_per_launch_isolated"), v.get("tflops_isolated"), v.get("gbs_isolated")))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench.err").read()[-3000:])
PY
