#!/bin/bash
# round 2, second pass: --set full captures of the non-GEMM-bound ECAPA kernels (592 crops, 16-bit mode)
mkdir -p gpurun_out
python tools/profile_forward.py 592 f16 2 > gpurun_out/pf.log 2>&1 || exit 1
cap() {  # name regex skip
  ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 \
      -o gpurun_out/r02b_$1 -f python tools/profile_forward.py 592 f16 2 > gpurun_out/ncu_r02b_$1.log 2>&1
}
cap chain res2net_chain 3
cap asp asp_fused 1
cap tdnn1 gemm_pw_kernel 10
cap fbank fbank_logmel 1
cap seapply se_apply 3
ls -la gpurun_out/r02b_*.ncu-rep
exit 0
