#!/bin/bash
# round 2, GPU call 24: ncu launch list (gpu__time_duration) of the bench command on the round's final binary
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident"
$B > gpurun_out/plain_full.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_final.csv $B > gpurun_out/ncu_full.log 2>&1
echo "ncu list rc=$?"
python tools/launch_summary.py gpurun_out/r02_launches_final.csv | tee gpurun_out/r02_launch_shares_final.txt | head -30
