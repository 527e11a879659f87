#!/bin/bash
# Round-2 first GPU visit: micro-benchmarks, self-test, parity tests, the new default bench line (configs[2] stage 1,
# B=32) and stage 2, ncu launch list of the B=32 step, ncu --set full of the LayerNorm-epilogue GEMM, comparators.
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt; lscpu | head -20 >> $O/smi.txt
cd future-object-detection_b200
timeout 120 ./microbench_fod > ../$O/microbench.log 2>&1; echo "microbench exit $?"
cat ../$O/microbench.log
timeout 300 ./selftest_fod 7 > ../$O/selftest.log 2>&1; echo "selftest exit $?" >> ../$O/selftest.log
tail -45 ../$O/selftest.log
cd ..
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_b32.json 2> $O/bench_b32.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_b32.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "gpu_launches", "clocks")})
    r = dict(d["roofline"]); print(json.dumps(r)[:2500])
    ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"])[:40]
    for k, v in ks: print("%-34s n=%4d ms/step %.4f iso_us %s tf %s gbs %s" % (k, v["launches"], v["ms_per_step"], v.get("us_per_launch_isolated"), v.get("tflops_isolated"), v.get("gbs_isolated")))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_b32.err").read()[-3000:])
PY
timeout 900 python bench.py --stage2 --steps 10 --warmup 3 > $O/bench_stage2.json 2> $O/bench_stage2.err; echo "stage2 exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_stage2.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "clocks")}); print(json.dumps(d["roofline"])[:2500])
    ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"])[:25]
    for k, v in ks: print("%-34s n=%4d ms/step %.4f iso_us %s tf %s gbs %s" % (k, v["launches"], v["ms_per_step"], v.get("us_per_launch_isolated"), v.get("tflops_isolated"), v.get("gbs_isolated")))
except Exception as e:
    print("stage2 parse failed", e); print(open("gpurun_out/bench_stage2.err").read()[-3000:])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_b32.csv \
  python bench.py --steps 1 --warmup 3 --graphs 0 --kernel-events 0 --no-cpu-baseline --profiler-range 1 > $O/ncu_bench.log 2>&1; echo "ncu list exit $?"
python scripts/summarize_launches.py $O/launches_b32.csv > $O/launches_b32.txt; head -75 $O/launches_b32.txt | cut -c1-170
gzip -f $O/launches_b32.csv
cd future-object-detection_b200
for spec in "lngemm_1024x256x256 5" "lngemm_1024x256x2048 16" "lngemm_22400x256x256 49"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 --launch-skip $2 --launch-count 1 -f -o ../$O/ncu_$1 ./selftest_fod 4 > ../$O/ncu_$1.log 2>&1; echo "ncu $1 exit $?"
done
cd ..
timeout 300 python scripts/sdpa_compare.py > $O/sdpa_compare.jsonl 2> $O/sdpa_compare.err; echo "sdpa exit $?"
cat $O/sdpa_compare.jsonl
timeout 300 python bench.py --impl reference --ref-device cuda --batch 8 --steps 10 --warmup 3 > $O/ref_gpu_f32.json 2> $O/ref_gpu_f32.err; echo "ref-gpu f32 exit $?"
timeout 300 python bench.py --impl reference --ref-device cuda --ref-autocast 1 --batch 8 --steps 10 --warmup 3 > $O/ref_gpu_bf16.json 2> $O/ref_gpu_bf16.err; echo "ref-gpu bf16 exit $?"
cat $O/ref_gpu_f32.json $O/ref_gpu_bf16.json
ls -la $O | head -50
