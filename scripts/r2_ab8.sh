#!/bin/bash
# A/B on ONE box: mid-step vs early num_boxes all-reduce, with/without the small tail bucket. usage: r2_ab8.sh N
N=$1
mkdir -p gpurun_out
O=gpurun_out
run() {  # tag early tail
  FOD_EARLY_NUM_BOXES=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 296$3$2 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --kernel-events 0 --dropout-off-leg 0 --tail-mb $3 > $O/ab_n${N}_$1.json 2> $O/ab_n${N}_$1.err; echo "$1 exit $?"
  python -c "
import json
d=json.load(open('$O/ab_n${N}_$1.json')); print('$1', {k: d[k] for k in ('n_gpus','value','ms_per_step')}, d['e2e']['value'], d['clocks'])"
}
run old 0 0
run new 1 4
run old2 0 0
run new2 1 4
