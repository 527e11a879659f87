#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py tests/test_optim_gpu.py -x -q -m gpu > gpurun_out/pytest_pw.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR|E  )|passed|failed" gpurun_out/pytest_pw.log | cut -c1-300 | head -20
for pw in 1 0; do
echo "FOD_PW_GEMM=$pw"; FOD_PW_GEMM=$pw bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 --dropout-off-leg 0 | grep -E "value|relu_bwd_add2|gemm_dx_3584|gemm_dw_.*3584|gemm_dx_8960|gemm_dw.*8960" | cut -c1-200
done
