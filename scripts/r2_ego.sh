#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_dropout_gpu.py tests/test_model_gpu.py -x -q -m gpu > $O/pytest_ego.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" $O/pytest_ego.log | cut -c1-300 | head -10
for m in 1 0; do
  FOD_EGO_BIAS_ROW=$m timeout 400 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --dropout-off-leg 0 2>/dev/null | tail -1 > $O/bench_ego_$m.json
  python - <<PY
import json
d=json.load(open("$O/bench_ego_$m.json"))
print("EGO_BIAS_ROW=$m", d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"])
PY
done
