#!/bin/bash
# Weak-scaling sweep on one box: N = 1, 2, 4, 8 ranks (torchrun, NCCL), each wrapped in its own timeout.
mkdir -p gpurun_out
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then
    timeout 240 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --kernel-events 0 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --kernel-events 0 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n exit $?"
  python -c "
import json
try:
    d = json.load(open('gpurun_out/scale_n$n.json')); print({k: d[k] for k in ('n_gpus', 'value', 'ms_per_step', 'e2e', 'loss_first_last_timed_step')})
except Exception as e:
    print('parse failed', e)
"
done
