#!/bin/bash
# A/B of two builds of libsvb in ONE call (box-to-box spread is 5 %): speakerverification_b200/libsvb_base.so vs libsvb.so
run() { timeout 300 python bench.py --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', round(d['value']), 'ms/step', round(d['ms_per_step'],2), 'mfa ms', round(d['roofline']['kernel_ms'],3), d['clocks']['sm_mhz'])"; }
timeout 1200 python -m pytest tests/test_tdnn_gpu.py tests/test_ecapa_gpu.py tests/test_fullsize_gpu.py tests/test_api_gpu.py tests/test_gemm_pw_gpu.py -m gpu -x -q 2>&1 | tail -3
for i in 1 2 3; do
echo -n "base: "; SVB_LIB=$PWD/speakerverification_b200/libsvb_base.so run
echo -n "new:  "; run
done
