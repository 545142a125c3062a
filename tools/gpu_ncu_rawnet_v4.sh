mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:rn_gemm_kernel --launch-skip 0 --launch-count 1 -o gpurun_out/r02_rn_sinc_v4 -f python tools/profile_rawnet.py 592 f16 1 > gpurun_out/ncu_r02_rn_sinc_v4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rn_gemm_kernel --launch-skip 1 --launch-count 2 -o gpurun_out/r02_rn_conv_v4 -f python tools/profile_rawnet.py 592 f16 1 > gpurun_out/ncu_r02_rn_conv_v4.log 2>&1
ls -la gpurun_out/*_v4.ncu-rep
