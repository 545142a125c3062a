#!/bin/bash
# A/B of a forward-path change: parity (layer, model, API incl. the bf16 list tolerance) + launch list + short bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_tdnn_gpu.py tests/test_ecapa_gpu.py tests/test_fullsize_gpu.py tests/test_api_gpu.py -m gpu -x -q 2>&1 | tail -4
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_fwd_q.csv python tools/profile_forward.py 592 bf16 2 > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/launches_fwd_q.csv | head -24
for i in 1 2; do
timeout 300 python bench.py --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', d['value'], 'ms/step', d['ms_per_step'], 'mfa TF', d['roofline']['achieved'], d['roofline']['kernel_ms'], d['clocks'])"
done
echo "--- SVB_PDL=0"
for i in 1 2; do
SVB_PDL=0 timeout 300 python bench.py --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', d['value'], 'ms/step', d['ms_per_step'], 'mfa TF', d['roofline']['achieved'], d['roofline']['kernel_ms'], d['clocks'])"
done
