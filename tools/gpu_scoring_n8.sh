#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --workload scoring --steps 20 > gpurun_out/r2_scoring_n8.json 2> gpurun_out/r2_scoring_n8.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_scoring_n8.json').read())
print('scoring N=8', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])
print(d.get('per_rank_ms_per_step'))
PY
