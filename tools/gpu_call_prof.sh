#!/bin/bash
# GPU call: ncu launch list at full size and --set full captures of the leaf hash and the NTT passes (final binary of the round)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_full.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_full.csv $B > gpurun_out/ncu_full.log 2>&1
echo "ncu list rc=$?"
$B --log-rows 18 > gpurun_out/plain_leaf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_linear_hash -s 1 -c 1 -o gpurun_out/prof_leaf $B --log-rows 18 > gpurun_out/ncu_leaf.log 2>&1
echo "ncu leaf rc=$?"
$B --cols 64 > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ntt_pass -s 6 -c 6 -o gpurun_out/prof_ntt $B --cols 64 > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
ls -la gpurun_out | tail -12
