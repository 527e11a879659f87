#!/bin/bash
# usage: r2_ncu_one.sh <name> <kernel regex> <launch skip> <selftest mask>   — one ncu --set full capture of a selftest launch
mkdir -p gpurun_out
cd future-object-detection_b200
timeout 120 ./selftest_fod $4 > ../gpurun_out/selftest_plain.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 -f -o ../gpurun_out/ncu_$1 ./selftest_fod $4 > ../gpurun_out/ncu_$1.log 2>&1
echo "ncu $1 exit $?"; tail -3 ../gpurun_out/ncu_$1.log
