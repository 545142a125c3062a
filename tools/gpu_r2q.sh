#!/bin/bash
SVB_FILTER_2SM=1 timeout 900 python -m pytest tests/test_scoring_gpu.py -m gpu -q 2>&1 | tail -2
for t in 0 1 0 1; do
SVB_FILTER_2SM=$t timeout 300 python bench.py --workload scoring --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('2SM=$t trials/s', d['value'], 'ms/step', d['ms_per_step'])"
done
