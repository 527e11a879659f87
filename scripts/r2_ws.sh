#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "gemm or ffn or linear" > $O/pytest_ws.log 2>&1; echo "pytest exit $?" >> $O/pytest_ws.log
grep -E "^(FAILED|ERROR|E  )|passed|failed|exit" $O/pytest_ws.log | cut -c1-300 | head -60
bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 "$@" | cut -c1-1200
