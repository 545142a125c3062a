#!/bin/bash
for br in 592 1184 592 1184 888; do
timeout 300 python bench.py --workload embed --no-e2e --steps 5 --batch-rows $br 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('batch_rows=$br emb/s', d['value'], 'ms/step', d['ms_per_step'], d['clocks']['sm_mhz'])"
done
