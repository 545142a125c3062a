#!/bin/bash
# the default bench line on N GPUs as the driver's scaling run launches it (usage: gpurun --gpus N -- bash tools/gpu_scale_all.sh N)
N=${1:-8}
mkdir -p gpurun_out
SECONDS=0
timeout 850 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err
echo "rc=$? wall ${SECONDS}s"
tail -3 gpurun_out/r2_scale_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_scale_n$N.json').read())
print('embed', d['value'], 'e2e', d['e2e']['value'], d.get('per_rank_ms_per_step'), d.get('strong_scaling'))
for k in ('eval','rawnet2','scoring'):
    print(k, d[k]['value'], d[k]['unit'], 'e2e', (d[k].get('e2e') or {}).get('value'))
print('api', d.get('api_distributed_check'))
PY
