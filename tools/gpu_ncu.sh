#!/bin/bash
# ncu evidence for the round: launch list + --set full captures of the top kernels at the bench launch size
B=${1:-592}
mkdir -p gpurun_out
python tools/profile_forward.py $B bf16 2 > gpurun_out/pf.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_fwd.csv \
    python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pw_kernel --launch-skip 14 --launch-count 1 \
    -o gpurun_out/full_mfa -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pw_kernel --launch-skip 9 --launch-count 1 \
    -o gpurun_out/full_tdnn2 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fbank_logmel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/full_fbank -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:res2net_chain --launch-skip 3 --launch-count 1 \
    -o gpurun_out/full_res2 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/full_block0 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f6.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:asp_fused --launch-skip 1 --launch-count 1 \
    -o gpurun_out/full_asp -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f7.log 2>&1
exit 0
