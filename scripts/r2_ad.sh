#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_dropout_gpu.py -q -m gpu -k "gemm or ln or ffn or block" 2>&1 | tail -1
for d in 1 0; do for c in "fwd_ln 256x2048" "fwd_ln 256x2048 drop" "fwd_ln_save 256x2048 drop"; do
echo -n "direct=$d "; FOD_GEMM_AUX_DIRECT=$d GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done; done
for d in 1 0; do echo "direct=$d"; FOD_GEMM_AUX_DIRECT=$d bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 --dropout-off-leg 0 | grep -E "value|gemm_fwd_ln_22400x256x2048 |gemm_fwd_ln_4096x256x2048|relu_bwd_add3_18" | cut -c1-160; done
