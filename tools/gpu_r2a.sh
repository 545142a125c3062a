#!/bin/bash
# round 2, call A: full GPU test suite (fp16 16-bit mode, K-packed block 0 on gemm_pw) + short bench + A/B of the block-0 layouts
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | grep -v "^frame #" | tail -150 > gpurun_out/r2a_pytest.txt
tail -5 gpurun_out/r2a_pytest.txt
for kp in 1 0; do
  SVB_BLOCK0_KPACK=$kp timeout 300 python bench.py --no-e2e --steps 3 2>gpurun_out/r2a_bench_kp$kp.err > gpurun_out/r2a_bench_kp$kp.json
  python -c "
import json,sys
d=json.loads(open('gpurun_out/r2a_bench_kp$kp.json').read()); print('kpack=$kp emb/s', d['value'], 'ms/step', d['ms_per_step'], 'mfa TF', d['roofline']['achieved'], d['clocks'])"
done
