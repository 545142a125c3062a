#!/bin/bash
# N-GPU bench lines exactly as the driver launches them: tools/gpu_scale.sh N [workloads...]
N=${1:-8}; shift
mkdir -p gpurun_out
for w in ${@:-embed}; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --workload $w > gpurun_out/bench_${w}_n$N.json 2> gpurun_out/bench_${w}_n$N.err
  tail -1 gpurun_out/bench_${w}_n$N.json | cut -c1-600
done
