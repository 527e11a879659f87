#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm or block" 2>&1 | tail -1
for c in "fwd_ln 256x2048" "fwd_ln_save 256x2048 drop" "dx 256x2048 residual" "dx 256x2048 plain"; do
GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done
bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 --dropout-off-leg 0 | grep -E "value|gemm_fwd_ln_22400x256x2048|gemm_dx_22400x256x2048|gemm_fwd_22400x256x2048|relu_bwd_add3_18" | cut -c1-200
