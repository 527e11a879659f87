#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "ln or block or linear or ffn" 2>&1 | tail -2
for c in "ln_bwd 22400" "ln_bwd 22400 drop" "ln_bwd 4096 drop" "ln_bwd 128 drop"; do
GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done
