#!/bin/bash
# round 2, call C: new tests (input_norm, hybrid, back-to-back staging, ragged AS-norm, pnorm) with printed deviations,
# launch list of one 592-crop forward for the K-packed block 0
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^frame #" | grep -E "dev|cos|passed|failed|Error|error|EER|FAIL|assert" | tail -80 > gpurun_out/r2c_pytest.txt
cat gpurun_out/r2c_pytest.txt | tail -40
for kp in 1 0; do
SVB_BLOCK0_KPACK=$kp ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2c_launches_kp$kp.csv python tools/profile_forward.py 592 bf16 2 > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/r2c_launches_kp$kp.csv | head -24
done
