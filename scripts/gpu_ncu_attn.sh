#!/bin/bash
mkdir -p gpurun_out
cd future-object-detection_b200
timeout 300 ncu --set full --clock-control none --import-source on -k regex:attn_fwd --launch-skip 58 --launch-count 1 -f -o ../gpurun_out/ncu_attn_fwd_v2_s1400 ./selftest_fod 2 > ../gpurun_out/ncu_attn_v2.log 2>&1
echo exit $?
