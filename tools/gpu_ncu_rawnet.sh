#!/bin/bash
# --set full captures of the three dominant RawNet2 GEMMs (592 utterances, 16-bit mode): sinc, layer1 conv1 / conv2 (W-resident)
mkdir -p gpurun_out
python tools/profile_rawnet.py 592 f16 1 > gpurun_out/prn.log 2>&1 || exit 1
cap() {
  ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 \
      -o gpurun_out/r02_rn_$1 -f python tools/profile_rawnet.py 592 f16 1 > gpurun_out/ncu_r02_rn_$1.log 2>&1
}
cap sinc rn_gemm_kernel 0
cap conv1 rn_gemm_kernel 1
cap conv2 rn_gemm_kernel 2
cap pool rn_pool_kernel 0
cap apply rn_apply_kernel 0
ls -la gpurun_out/r02_rn_*.ncu-rep
exit 0
