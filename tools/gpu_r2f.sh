#!/bin/bash
# round 2, call F (2 GPUs): the default bench line at N = 2 (sub-lines, strong scaling, distributed API check)
mkdir -p gpurun_out
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2f_n2.json 2> gpurun_out/r2f_n2.err
echo "rc=$? wall ${SECONDS}s"
tail -5 gpurun_out/r2f_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2f_n2.json').read())
print('embed', d['value'], 'e2e', d['e2e']['value'], d.get('per_rank_ms_per_step'), d.get('strong_scaling'))
for k in ('eval','rawnet2','scoring'):
    print(k, d[k]['value'], d[k]['unit'], 'e2e', (d[k].get('e2e') or {}).get('value'))
print('api', d.get('api_distributed_check'))
PY
