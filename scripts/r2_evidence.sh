#!/bin/bash
# Round-2 evidence in one call: GPU test tier + smoke, default bench (with cpu_baseline), reference arm, stage 2, op bench,
# micro-benchmarks, sdpa comparison, eager GPU reference, ncu launch lists (p = 0.1 and p = 0).  ncu --set full: r2_ncu_ops.sh.
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/smi.txt 2>&1
bash scripts/r2_full.sh
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"; cut -c1-600 $O/bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"; cut -c1-400 $O/bench_ref.json
timeout 900 python bench.py --stage2 --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_stage2.json 2> $O/bench_stage2.err; echo "stage2 exit $?"; cut -c1-400 $O/bench_stage2.json
timeout 300 python scripts/gemm_bench.py > $O/op_bench.jsonl 2> $O/op_bench.err; echo "opbench exit $?"; cat $O/op_bench.jsonl
(cd future-object-detection_b200 && timeout 120 ./microbench_fod) > $O/microbench.log 2>&1; cat $O/microbench.log
timeout 600 python scripts/sdpa_compare.py > $O/sdpa_compare.jsonl 2> $O/sdpa_compare.err; echo "sdpa exit $?"; cat $O/sdpa_compare.jsonl
timeout 600 python bench.py --impl reference --ref-device cuda --steps 5 --warmup 3 > $O/ref_gpu_f32.json 2> $O/ref_gpu_f32.err; cut -c1-300 $O/ref_gpu_f32.json
timeout 600 python bench.py --impl reference --ref-device cuda --ref-autocast 1 --steps 5 --warmup 3 > $O/ref_gpu_bf16.json 2> $O/ref_gpu_bf16.err; cut -c1-300 $O/ref_gpu_bf16.json
(cd future-object-detection_b200 && timeout 300 ./selftest_fod) > $O/selftest.log 2>&1; echo "selftest exit $?"; tail -5 $O/selftest.log | cut -c1-200
bash scripts/r2_launches.sh b32_p01 | head -30
bash scripts/r2_launches.sh b32_p0 --dropout 0 | head -12
