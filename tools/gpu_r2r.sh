#!/bin/bash
timeout 900 python -m pytest tests/test_scoring_gpu.py -m gpu -q -s 2>&1 | grep -E "fused|passed|failed" | tail -4
for t in 1 2; do
timeout 300 python bench.py --workload scoring --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('trials/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2r_launches_scoring.csv python tools/profile_scoring.py 65536 > /dev/null 2>&1
python tools/launch_agg.py gpurun_out/r2r_launches_scoring.csv 8
