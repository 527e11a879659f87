#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/gemm_bench.py > gpurun_out/gemm_bench.jsonl 2> gpurun_out/gemm_bench.err; echo "exit $?"; cat gpurun_out/gemm_bench.jsonl; tail -3 gpurun_out/gemm_bench.err
FOD_GEMM_WS=0 timeout 300 python scripts/gemm_bench.py > gpurun_out/gemm_bench_ws0.jsonl 2>> gpurun_out/gemm_bench.err; echo "exit $?"; cat gpurun_out/gemm_bench_ws0.jsonl
