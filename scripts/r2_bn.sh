#!/bin/bash
for bn in 0 128 64; do for c in "dx 256x2048 residual" "dx 256x2048 plain" "fwd 256x256 bias"; do
echo -n "BN=$bn "; FOD_GEMM_BN=$bn GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py 2>&1 | tail -1
done; done
