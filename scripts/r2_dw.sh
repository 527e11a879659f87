#!/bin/bash
for m in 22400 4096 44800; do for c in "dw 2048x256" "dw 256x2048" "dw 768x256" "dw 256x256"; do
echo -n "M=$m "; GEMM_CASE="$c" timeout 120 python scripts/gemm_bench.py $m 2>&1 | tail -1
done; done
bash scripts/r2_full_bench.sh | grep -E "passed|failed|exit|value|dropout_off" | cut -c1-300
