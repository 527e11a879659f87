#!/bin/bash
# ncu launch list (gpu__time_duration per launch, cold-cache/serialised) of one eager step: usage r2_launches.sh <tag> [bench args]
mkdir -p gpurun_out
O=gpurun_out
T=$1; shift
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/launches_$T.csv \
  python bench.py --steps 1 --warmup 3 --graphs 0 --kernel-events 0 --no-cpu-baseline --dropout-off-leg 0 --profiler-range 1 "$@" > $O/ncu_bench_$T.log 2>&1; echo "ncu list exit $?"
python scripts/summarize_launches.py $O/launches_$T.csv > $O/launches_$T.txt; head -60 $O/launches_$T.txt | cut -c1-170
gzip -f $O/launches_$T.csv
