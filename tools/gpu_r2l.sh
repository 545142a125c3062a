#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_scoring_gpu.py tests/test_loader_gpu.py tests/test_rawnet_gpu.py -m gpu -q 2>&1 | tail -3
timeout 300 python bench.py --workload scoring --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('scoring trials/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
timeout 300 python bench.py --workload eval --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('eval utt/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
