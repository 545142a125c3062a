#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_scoring_gpu.py -m gpu -q 2>&1 | tail -2
ncu --set full --clock-control none -k regex:gemm_pw_kernel --launch-skip 3 --launch-count 1 -o gpurun_out/r02_filter32k -f python tools/profile_scoring.py 65536 > gpurun_out/ncu_r02_filter32k.log 2>&1
ncu -i gpurun_out/r02_filter32k.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h,u,v=rows[0],rows[1],rows[2]; ix={n:i for i,n in enumerate(h)}
for k in ('Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed','lts__t_sector_hit_rate.pct','launch__grid_size'):
    print(k, v[ix[k]], u[ix[k]])"
for ch in 32768 65536; do
SVB_TOPK_CHUNK=$ch timeout 300 python bench.py --workload scoring --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunk=$ch trials/s', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])"
done
