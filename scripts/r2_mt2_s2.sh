#!/bin/bash
# stage-2 (896x1600, 44 800 token rows = 350 tiles) A/B of the two-tile GEMM path
mkdir -p gpurun_out
O=gpurun_out
for m in 2 0; do
  FOD_GEMM_MT2=$m timeout 900 python bench.py --stage2 --steps 15 --warmup 4 --no-cpu-baseline --dropout-off-leg 0 2>/dev/null | tail -1 > $O/bench_s2_mt2_$m.json
  python - <<PY
import json
d=json.load(open("$O/bench_s2_mt2_$m.json"))
print("stage2 MT2=$m", d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"])
PY
done
