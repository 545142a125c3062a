#!/bin/bash
# round 2, call G: full GPU suite (RawNet2 heads, fbank TMA staging, se_apply), launch list, short bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^frame #" | grep -E "RawNet2 s|RawNet2 a|passed|failed|Error|error|assert|FAIL" | tail -30
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2g_launches.csv python tools/profile_forward.py 592 f16 2 > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/r2g_launches.csv | head -20
timeout 300 python bench.py --workload embed --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('emb/s', d['value'], 'ms/step', d['ms_per_step'], d['clocks'])"
