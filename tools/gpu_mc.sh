#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tdnn_gpu.py tests/test_ecapa_gpu.py -m gpu -x -q 2>&1 | tail -3
for mc in 1 0; do
echo "SVB_PW_MC=$mc"
SVB_PW_MC=$mc timeout 300 python bench.py --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', d['value'], 'ms/step', d['ms_per_step'], 'mfa TF', d['roofline']['achieved'], d['roofline']['kernel_ms'])"
done
SVB_PW_MC=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_fwd_mc.csv python tools/profile_forward.py 592 bf16 2 > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_fwd_mc.csv | head -8
