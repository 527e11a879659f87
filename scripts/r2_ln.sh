#!/bin/bash
mkdir -p gpurun_out
(cd future-object-detection_b200 && timeout 300 ./selftest_fod 255 > ../gpurun_out/selftest_all.log 2>&1; echo "selftest exit $?"; tail -2 ../gpurun_out/selftest_all.log | cut -c1-150)
timeout 300 python scripts/gemm_bench.py > gpurun_out/op_bench.jsonl 2> gpurun_out/op_bench.err; cat gpurun_out/op_bench.jsonl
bash scripts/r2_full_bench.sh | grep -E "passed|failed|exit|value|dropout_off" | cut -c1-300
