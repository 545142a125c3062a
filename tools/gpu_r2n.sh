#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload scoring --steps 20 > gpurun_out/r2n_scoring_n$N.json 2> gpurun_out/r2n_scoring_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2n_scoring_n$N.json').read())
print('scoring N=$N', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'])
print(d.get('per_rank_ms_per_step'))
print(d.get('clocks'))
PY
nvidia-smi --query-gpu=index,clocks.sm,power.draw,power.limit,temperature.gpu --format=csv
nproc
