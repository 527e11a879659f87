#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log | cut -c1-300
for m in "1 1" "0 1" "1 0" "1 1" "0 1"; do
  set -- $m
  FOD_ATTN_BITS=$1 FOD_ATTN_BITS_PREGEN=$2 timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --dropout-off-leg 0 2>$O/bench_bits_err.log | tail -1 > $O/bench_bits_$1$2.json
  python - <<PY
import json
d=json.load(open("$O/bench_bits_$1$2.json"))
print("ATTN_BITS=$1 PREGEN=$2", d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"], d["gpu_launches"])
PY
done
