#!/bin/bash
bash scripts/r2_full.sh
mkdir -p gpurun_out
timeout 900 python bench.py --stage2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_stage2.json 2> gpurun_out/bench_stage2.err; echo "stage2 exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_stage2.json'))
print({k: d[k] for k in ("value","ms_per_step","clocks")}, d['e2e']['value'], d.get('dropout_off'))
r=dict(d['roofline']); att=r.pop('attention'); print(r['kernel'], r['avg_launch_us'], r['frac'], [(x['kernel'], round(x['ms_per_step'],3)) for x in r['runners_up'][:4]])
print({k:(v['kernel'],round(v['avg_launch_us'],1),round(v['tflops'],1)) for k,v in att.items()})
PY
bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 | grep -E "value" | cut -c1-200
