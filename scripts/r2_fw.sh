#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "attention" > gpurun_out/pytest_fw.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_fw.log | cut -c1-300 | head -30
for c in "attn_fwd 64x350" "attn_fwd 64x350 drop" "attn_fwd 32x1400"; do
GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done
(cd future-object-detection_b200 && timeout 120 ./selftest_fod 2 2>&1 | grep -E "attn|SELF" | cut -c1-160)
