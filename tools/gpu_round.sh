#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench lines, ncu launch list and --set full captures (each after its plain run).
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench_embed.json 2> gpurun_out/bench_embed.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --workload eval > gpurun_out/bench_eval.json 2> gpurun_out/bench_eval.err
python bench.py --workload rawnet2 > gpurun_out/bench_rawnet2.json 2> gpurun_out/bench_rawnet2.err
python bench.py --workload scoring > gpurun_out/bench_scoring.json 2> gpurun_out/bench_scoring.err
B=${1:-592}
python tools/profile_forward.py $B bf16 2 > gpurun_out/pf.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_fwd.csv \
    python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pw_kernel --launch-skip 14 --launch-count 1 \
    -o gpurun_out/full_mfa -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_pw_kernel --launch-skip 9 --launch-count 1 \
    -o gpurun_out/full_tdnn2 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f5.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fbank_logmel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/full_fbank -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:res2net_chain --launch-skip 3 --launch-count 1 \
    -o gpurun_out/full_res2 -f python tools/profile_forward.py $B bf16 2 > gpurun_out/ncu_f3.log 2>&1
exit 0
