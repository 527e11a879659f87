#!/bin/bash
# two-tile-per-CTA GEMM path: tests, then the op bench with the path on / off / extended to short contractions
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm" > gpurun_out/pytest_mt2.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_mt2.log | cut -c1-300 | head -40
M="fwd_ln,dx 256x2048,fwd 256x256"
echo "== MT2 on (default)"; GEMM_MATCH="$M" timeout 300 python scripts/gemm_bench.py 2>&1 | grep case
echo "== MT2 off"; FOD_GEMM_MT2=0 GEMM_MATCH="$M" timeout 300 python scripts/gemm_bench.py 2>&1 | grep case
echo "== MT2 incl. short K"; FOD_GEMM_MT2=2 GEMM_MATCH="$M" timeout 300 python scripts/gemm_bench.py 2>&1 | grep case
