#!/bin/bash
# round 2, call D: rawnet f16 tests, restructured bench (small, then full default)
mkdir -p gpurun_out
echo "--- small bench"
timeout 600 python bench.py --steps 3 --utts 1184 --eval-utts 200 --eval-trials 1000 --score-utts 20000 --score-trials 10000 > gpurun_out/r2d_small.json 2> gpurun_out/r2d_small.err || tail -30 gpurun_out/r2d_small.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2d_small.json').read())
    print('small ok', d['value'], list(d.keys()))
except Exception as e:
    print('small failed', e)
PY
echo "--- full default bench"
SECONDS=0; timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2d_full.json 2> gpurun_out/r2d_full.err
echo "full bench wall seconds: $SECONDS"; tail -3 gpurun_out/r2d_full.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2d_full.json').read())
print('embed', d['value'], 'e2e', d['e2e']['value'], 'step frac', d['step_tensor']['frac'], d['clocks'])
for k in ('eval','rawnet2','scoring'):
    print(k, d[k]['value'], d[k]['unit'], 'e2e', (d[k].get('e2e') or {}).get('value'), 'roof', d[k]['roofline']['frac'], 'cpu', (d[k].get('cpu_baseline') or {}).get('value'))
print('lib', d.get('gpu_library_baseline'))
print('cpu', d.get('cpu_baseline'))
PY
