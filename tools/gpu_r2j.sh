#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rawnet_gpu.py -m gpu -q -s 2>&1 | grep -E "RawNet2 f16|front end f16|passed|failed|Error" | tail -6
for w in 1 0; do
SVB_RN_WRES=$w timeout 300 python bench.py --workload rawnet2 --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('WRES=$w rawnet2 emb/s', d['value'], 'ms/step', d['ms_per_step'])"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2j_launches_rawnet.csv python tools/profile_rawnet.py 256 f16 1 > /dev/null 2>&1
python tools/launch_agg.py gpurun_out/r2j_launches_rawnet.csv 10
