#!/bin/bash
# round 2: the full GPU suite + the default bench line (N = 1) exactly as the driver runs it + the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4
SECONDS=0
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
echo "bench rc=$? wall ${SECONDS}s"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n1.json').read())
print('embed', d['value'], 'e2e', d['e2e']['value'], 'step frac', d['step_tensor']['frac'], 'mfa', d['roofline']['achieved'], d['clocks'])
for k in ('eval','rawnet2','scoring'):
    print(k, d[k]['value'], d[k]['unit'], 'e2e', (d[k].get('e2e') or {}).get('value'), 'roof', d[k]['roofline']['frac'], 'cpu', (d[k].get('cpu_baseline') or {}).get('value'))
print('lib', json.dumps(d.get('gpu_library_baseline')))
print('cpu', d.get('cpu_baseline'))
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | cut -c1-300
