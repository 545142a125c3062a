#!/bin/bash
# round 2: --set full captures of every kernel class of the forward (592 crops, 16-bit mode) and of the fused AS-norm path
B=${1:-592}
mkdir -p gpurun_out
python tools/profile_forward.py $B f16 2 > gpurun_out/pf.log 2>&1 || exit 1
cap() {  # name regex skip
  ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 \
      -o gpurun_out/r02_$1 -f python tools/profile_forward.py $B f16 2 > gpurun_out/ncu_r02_$1.log 2>&1
}
cap block0 gemm_pw_kernel 9
cap tdnn1 gemm_pw_kernel 10
cap tdnn2 gemm_pw_kernel 11
cap mfa gemm_pw_kernel 16
cap chain res2net_chain 3
cap seapply se_apply 3
cap asp asp_fused 1
cap fbank fbank_logmel 1
ncu --set full --clock-control none --import-source on -k regex:gemm_pw_kernel --launch-skip 3 --launch-count 1 \
    -o gpurun_out/r02_filter -f python tools/profile_scoring.py 16384 > gpurun_out/ncu_r02_filter.log 2>&1
ncu --set full --clock-control none -k regex:cand_select --launch-skip 2 --launch-count 1 \
    -o gpurun_out/r02_candsel -f python tools/profile_scoring.py 16384 > gpurun_out/ncu_r02_candsel.log 2>&1
ls -la gpurun_out/*.ncu-rep
exit 0
