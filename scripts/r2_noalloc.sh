#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_zzz_no_alloc.py -q -m gpu > $O/pytest_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -4 $O/pytest_noalloc_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "full exit $?"; tail -4 $O/pytest_gpu.log | cut -c1-300
bash scripts/r2_ncu_ops.sh
