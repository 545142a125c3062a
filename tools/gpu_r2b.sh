#!/bin/bash
mkdir -p gpurun_out
for t in tests/test_gemm_pw_gpu.py tests/test_tdnn_gpu.py tests/test_ecapa_gpu.py; do
  echo "=== $t" >> gpurun_out/r2b.txt
  CUDA_LAUNCH_BLOCKING=1 timeout 600 python -m pytest $t -m gpu -q -x 2>&1 | grep -v "^frame #" | head -150 >> gpurun_out/r2b.txt
done
