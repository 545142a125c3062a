#!/bin/bash
# round 2, call E: fused AS-norm path: parity tests, A/B against the store-and-select path, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_scoring_gpu.py tests/test_api_gpu.py -m gpu -q -s 2>&1 | grep -v "^frame #" | grep -E "fused|passed|failed|Error|error|assert|dev" | tail -30
for f in 1 0; do
  SVB_TOPK_FUSED=$f timeout 300 python bench.py --workload scoring --steps 10 2>gpurun_out/r2e_sc$f.err > gpurun_out/r2e_scoring_fused$f.json
  python -c "
import json
d=json.loads(open('gpurun_out/r2e_scoring_fused$f.json').read()); print('fused=$f trials/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'])" || tail -5 gpurun_out/r2e_sc$f.err
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2e_launches_scoring.csv python tools/profile_scoring.py 16384 > /dev/null 2>&1
python tools/launch_agg.py gpurun_out/r2e_launches_scoring.csv
