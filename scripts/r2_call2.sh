#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
FOD_TEST_DUMP=$O timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
for cb in 1 0; do
timeout 900 python bench.py --steps 20 --warmup 5 --cudnn-benchmark $cb > $O/bench_b32_cb$cb.json 2> $O/bench_b32_cb$cb.err; echo "bench cb=$cb exit $?"
python - $cb <<'PY'
import json, sys
cb = sys.argv[1]
try:
    d = json.load(open("gpurun_out/bench_b32_cb%s.json" % cb))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "gpu_launches", "clocks")})
    r = dict(d["roofline"]); print(json.dumps(r)[:3000])
    ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"])[:45]
    for k, v in ks: print("%-34s n=%4d ms/step %.4f iso_us %s tf %s gbs %s" % (k, v["launches"], v["ms_per_step"], v.get("us_per_launch_isolated"), v.get("tflops_isolated"), v.get("gbs_isolated")))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_b32_cb%s.err" % cb).read()[-3000:])
PY
done
timeout 900 python bench.py --stage2 --steps 10 --warmup 3 > $O/bench_stage2.json 2> $O/bench_stage2.err; echo "stage2 exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_stage2.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks")}); print(json.dumps(d["roofline"])[:3000])
    ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"])[:25]
    for k, v in ks: print("%-34s n=%4d ms/step %.4f iso_us %s tf %s gbs %s" % (k, v["launches"], v["ms_per_step"], v.get("us_per_launch_isolated"), v.get("tflops_isolated"), v.get("gbs_isolated")))
except Exception as e:
    print("stage2 parse failed", e); print(open("gpurun_out/bench_stage2.err").read()[-3000:])
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"; cat $O/bench_ref.json
