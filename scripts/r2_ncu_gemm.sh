#!/bin/bash
# usage: r2_ncu_gemm.sh "<case name>" <tag> [kernel regex]
mkdir -p gpurun_out
K=${3:-gemm_ws}
GEMM_CASE="$1" GEMM_EAGER=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o gpurun_out/ncu_$2 -f python scripts/gemm_bench.py > gpurun_out/ncu_$2.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_$2.log
