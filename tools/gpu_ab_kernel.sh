#!/bin/bash
# A/B one kernel's duration (ncu launch list) between libsvb_base.so and libsvb.so, then parity + bench A/B.
# usage: tools/gpu_ab_kernel.sh <kernel-regex>
K=${1:-fbank_logmel}
mkdir -p gpurun_out
for lib in libsvb_base.so libsvb.so; do
  SVB_LIB=$PWD/speakerverification_b200/$lib ncu --metrics gpu__time_duration.sum --clock-control none -k regex:$K -c 8 --csv \
      --log-file gpurun_out/ab_$lib.csv python tools/profile_forward.py 592 bf16 2 > /dev/null 2>&1
  echo -n "$lib $K us: "; python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/ab_$lib.csv')))
h=[i for i,r in enumerate(rows) if 'Metric Value' in r][0]
vi=rows[h].index('Metric Value')
print([round(float(r[vi].replace(',',''))/1000,1) for r in rows[h+2:] if len(r)>vi])
PY
done
bash tools/gpu_ab3.sh
