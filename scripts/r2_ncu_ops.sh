#!/bin/bash
# ncu --set full of one launch of each dominant kernel at its B = 32 benchmark shape (scripts/gemm_bench.py cases)
mkdir -p gpurun_out
cap() {  # tag, kernel regex, case name, skip, count
  GEMM_CASE="$3" GEMM_EAGER=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:$2 -s ${4:-1} -c ${5:-1} -f -o gpurun_out/ncu_r2_$1 python scripts/gemm_bench.py > gpurun_out/ncu_r2_$1.log 2>&1; echo "$1 exit $?"
}
cap ffn2ln gemm_bf16 "fwd_ln_save 256x2048 drop"
cap ffn1_ws gemm_ws "fwd 2048x256 bias+relu+drop"
cap dh_ws gemm_ws "dx 2048x256 mask+colsum"
cap dx_k2048 gemm_bf16 "dx 256x2048 residual"
cap attn_fwd_350 attn_fwd "attn_fwd 64x350 drop"
cap attn_bwd_350 attn_bwd_fused "attn_bwd 64x350 drop"
cap attn_fwd_1400 attn_fwd "attn_fwd 32x1400"
cap attn_bwd_1400 "attn_delta\|attn_bwd" "attn_bwd 32x1400" 3 3
cap relu_bwd_add3 relu_bwd_add "relu_bwd_add3 183500800"
ls -la gpurun_out/ncu_r2_*.ncu-rep
