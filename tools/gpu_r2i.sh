#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rawnet_gpu.py -m gpu -q 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches_rawnet.csv python tools/profile_rawnet.py 256 f16 2 > /dev/null 2>&1
python tools/launch_agg.py gpurun_out/r2i_launches_rawnet.csv 16
timeout 300 python bench.py --workload rawnet2 --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('rawnet2 emb/s', d['value'], 'ms/step', d['ms_per_step'])"
