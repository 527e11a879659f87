#!/bin/bash
# ncu launch list (gpu__time_duration per launch) of ONE eager bench step, captured between cudaProfilerStart/Stop.
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 3 --graphs 0 --kernel-events 0 --no-cpu-baseline --profiler-range 1 > gpurun_out/ncu_bench.log 2>&1; echo "ncu exit $?"
python scripts/summarize_launches.py gpurun_out/launches.csv > gpurun_out/launches_summary.txt; head -50 gpurun_out/launches_summary.txt | cut -c1-160
