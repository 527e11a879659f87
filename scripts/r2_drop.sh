#!/bin/bash
# dropout tests (all failures shown) + bench with dropout on/off
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_dropout_gpu.py -q -m gpu > $O/pytest_drop.log 2>&1; echo "pytest exit $?" >> $O/pytest_drop.log
grep -E "^(FAILED|ERROR|E  )|passed|failed|exit" $O/pytest_drop.log | cut -c1-400 | head -80
bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 "$@" | cut -c1-1500
