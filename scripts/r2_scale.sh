#!/bin/bash
# usage: r2_scale.sh N — the default bench line at N GPUs (torchrun), saved as gpurun_out/scale_final_nN.json
N=$1
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2967$N bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --dropout-off-leg 0 > gpurun_out/scale_final_n$N.json 2> gpurun_out/scale_final_n$N.err; echo "exit $?"
python -c "
import json; d=json.load(open('gpurun_out/scale_final_n$N.json')); print({k: d[k] for k in ('n_gpus','value','ms_per_step','clocks')}, d['e2e']['value'], d['e2e']['synchronous']['value'])"
