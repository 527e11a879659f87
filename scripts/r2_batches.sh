#!/bin/bash
# one-GPU step time at the per-GPU batches of the reference-faithful strong-scaling split (global 32 over 2 / 4 / 8 ranks)
mkdir -p gpurun_out
for b in 16 8 4; do
  for p in 0.1 0; do
    timeout 300 python bench.py --batch $b --dropout $p --steps 30 --warmup 5 --no-cpu-baseline --dropout-off-leg 0 2>/dev/null | tail -1 > gpurun_out/bench_b${b}_p${p}.json
    python - <<PY
import json
d=json.load(open("gpurun_out/bench_b${b}_p${p}.json"))
print("batch $b dropout $p:", round(d["ms_per_step"],3), "ms", round(d["value"],1), "clips/s", d["clocks"]["sm_mhz"], "MHz", d["gpu_launches"]//30, "launches/step")
PY
  done
done
