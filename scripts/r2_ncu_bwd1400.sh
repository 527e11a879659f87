#!/bin/bash
mkdir -p gpurun_out
GEMM_CASE="attn_bwd 32x1400" GEMM_EAGER=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_fused -s 1 -c 1 -f -o gpurun_out/ncu_r2_attn_bwd_1400 python scripts/gemm_bench.py > gpurun_out/ncu_r2_attn_bwd_1400.log 2>&1; echo "exit $?"; tail -2 gpurun_out/ncu_r2_attn_bwd_1400.log
