#!/bin/bash
# A/B a gemm_pw change: parity of the layer tests + the stand-alone timing + a short in-step bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_pw_gpu.py tests/test_tdnn_gpu.py tests/test_ecapa_gpu.py tests/test_fullsize_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/pw_alone.py 2>&1 | tail -8
for i in 1 2; do
timeout 300 python bench.py --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', d['value'], 'ms/step', d['ms_per_step'], 'mfa TF', d['roofline']['achieved'], d['roofline']['kernel_ms'], d['clocks'])"
done
