#!/bin/bash
# usage: r2_dist.sh N [bucket MB] — timeline of one N-GPU replay + the bench line at N
N=$1; BK=${2:-32}
mkdir -p gpurun_out
O=gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 scripts/profile_graph_dist.py 32 $BK 0.1 > $O/timeline_n${N}_b${BK}.log 2> $O/timeline_n${N}.err; echo "timeline exit $?"
cat $O/timeline_n${N}_b${BK}.log | cut -c1-200 | head -60; tail -3 $O/timeline_n${N}.err | cut -c1-300
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --bucket-mb $BK > $O/bench_n${N}_b${BK}.json 2> $O/bench_n${N}.err; echo "bench exit $?"
python -c "
import json
d=json.load(open('$O/bench_n${N}_b${BK}.json')); print({k: d[k] for k in ('n_gpus','value','ms_per_step','e2e','dropout_off','clocks')})"
