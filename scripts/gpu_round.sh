#!/bin/bash
# One GPU-box visit: parity tests, bench, ncu launch list of the same command, torch-profiler table. Outputs -> gpurun_out/
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
lscpu | head -20 >> gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 3000 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
cat gpurun_out/bench_ref.json
timeout 600 python scripts/profile_step.py > gpurun_out/profile_step.log 2>&1; echo "profile_step exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 3 --graphs 0 --kernel-events 0 --no-cpu-baseline --profiler-range 1 > gpurun_out/ncu_bench.log 2>&1; echo "ncu exit $?"
wc -l gpurun_out/launches.csv
# stock comparator for the encoder attention and the eager torch port on the same GPU
timeout 300 python scripts/sdpa_compare.py > gpurun_out/sdpa_compare.jsonl 2> gpurun_out/sdpa_compare.err; echo "sdpa exit $?"
cat gpurun_out/sdpa_compare.jsonl
timeout 200 python bench.py --impl reference --ref-device cuda --steps 10 --warmup 3 > gpurun_out/ref_gpu_f32.json 2> gpurun_out/ref_gpu_f32.err; echo "ref-gpu exit $?"
