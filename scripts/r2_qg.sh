#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "attention" > gpurun_out/pytest_qg.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_qg.log | cut -c1-300 | head -40
for c in "attn_bwd 32x1400" "attn_bwd 64x350" "attn_bwd 64x350 drop"; do
GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done
timeout 300 python scripts/sdpa_compare.py 2>/dev/null | cut -c1-260
