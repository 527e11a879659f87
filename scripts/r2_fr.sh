#!/bin/bash
timeout 600 python -m pytest tests/test_dropout_gpu.py -q -m gpu 2>&1 | grep -E "^E |passed|failed" | head
bash scripts/r2_full_bench.sh | grep -E "passed|failed|exit|value|dropout_off" | cut -c1-260
