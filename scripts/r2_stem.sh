#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "stem or frames" > gpurun_out/pytest_stem.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_stem.log | cut -c1-300 | head -20
GEMM_MATCH="stem,frames" timeout 300 python scripts/gemm_bench.py 2>&1 | grep "case\|Error\|error" | head
