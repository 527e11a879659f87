#!/bin/bash
# whole-step A/B of the fused stem convolution + max-pool (FOD_STEM_POOL) on one box, after the full GPU tier
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log | cut -c1-300
GEMM_MATCH="stem" timeout 300 python scripts/gemm_bench.py 2>&1 | grep "case" > $O/op_stem.jsonl; cat $O/op_stem.jsonl
for m in 1 0 1 0; do
  FOD_STEM_POOL=$m timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --dropout-off-leg 0 2>/dev/null | tail -1 > $O/bench_pool_$m.json
  python - <<PY
import json
d=json.load(open("$O/bench_pool_$m.json"))
print("STEM_POOL=$m", d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"])
PY
done
