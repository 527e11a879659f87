#!/bin/bash
# End-of-round evidence (v7): smoke + self-test, default bench, stage 2, op bench, ncu launch list at p = 0.1
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/smi.txt 2>&1
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?" >> $O/smoke.log; tail -3 $O/smoke.log | cut -c1-400
(cd future-object-detection_b200 && timeout 300 ./selftest_fod) > $O/selftest.log 2>&1; echo "selftest exit $?" >> $O/selftest.log; tail -3 $O/selftest.log | cut -c1-200
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"; cut -c1-400 $O/bench_default.json
timeout 900 python bench.py --stage2 --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_stage2.json 2> $O/bench_stage2.err; echo "stage2 exit $?"; cut -c1-300 $O/bench_stage2.json
timeout 400 python scripts/gemm_bench.py > $O/op_bench.jsonl 2> $O/op_bench.err; echo "opbench exit $?"; cat $O/op_bench.jsonl | cut -c1-120
bash scripts/r2_launches.sh b32_p01 | head -40
