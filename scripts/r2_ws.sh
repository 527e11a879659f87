#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "gemm or linear or ffn" > gpurun_out/pytest_ws.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_ws.log | cut -c1-300 | head -20
GEMM_MATCH="2048x256,768x256" timeout 300 python scripts/gemm_bench.py 2>&1 | grep case
