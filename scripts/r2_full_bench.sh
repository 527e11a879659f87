#!/bin/bash
bash scripts/r2_full.sh
bash scripts/r2_bench.sh --no-cpu-baseline --steps 10 "$@" | cut -c1-1500
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print('dropout_off', d.get('dropout_off'))"
