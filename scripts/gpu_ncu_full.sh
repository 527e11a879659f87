#!/bin/bash
# ncu --set full captures of the dominant kernels at their benchmark shapes (one launch each).
mkdir -p gpurun_out
cd future-object-detection_b200
cap() {  # name, kernel regex, skip, selftest arg
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 -f -o ../gpurun_out/$1 ./selftest_fod $4 > ../gpurun_out/$1.log 2>&1; echo "$1 exit $?"
}
if [ "$1" != "attn" ]; then
cap ncu_gemm_ffn1_5600x2048x256 gemm_bf16 25 1
cap ncu_gemm_ffn2ln_5600x256x2048 gemm_bf16 36 1
fi
cap ncu_attn_fwd_b16_s350 attn_fwd 47 2
cap ncu_attn_fwd_b16_s1400 attn_fwd 58 2
cd ..
# attention backward (delta + dq + dkdv kernels of the third call)
timeout 400 ncu --set full --clock-control none --import-source on -k regex:attn_delta\|attn_bwd --launch-skip 6 --launch-count 3 -f -o gpurun_out/ncu_attn_bwd_b16_s350 python scripts/ncu_attn_bwd.py > gpurun_out/ncu_attn_bwd.log 2>&1; echo "attn_bwd exit $?"
ls -la gpurun_out/*.ncu-rep
