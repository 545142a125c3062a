#!/bin/bash
for ch in 8192 16384 32768 8192 16384; do
SVB_TOPK_CHUNK=$ch timeout 300 python bench.py --workload scoring --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunk=$ch trials/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])"
done
