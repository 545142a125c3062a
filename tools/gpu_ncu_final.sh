#!/bin/bash
# --set full captures of every kernel class of the final forward (592 crops, bf16), one launch each
B=${1:-592}
mkdir -p gpurun_out
python tools/profile_forward.py $B bf16 2 > gpurun_out/pf.log 2>&1 || exit 1
cap() {  # name regex skip
  ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 \
      -o gpurun_out/final_$1 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_final_$1.log 2>&1
}
cap block0 gemm_pw_kernel 9
cap tdnn1 gemm_pw_kernel 10
cap tdnn2 gemm_pw_kernel 11
cap mfa gemm_pw_kernel 16
cap asptdnn gemm_pw_kernel 17
cap chain res2net_chain 3
cap seapply se_apply 3
cap asp asp_fused 1
cap fbank fbank_logmel 1
exit 0
