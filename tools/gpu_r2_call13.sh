#!/bin/bash
# round 2, GPU call 13: ncu source-level profile of the NTT passes (where do the 150 instructions per element go?)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify --cols 64"
$B > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ntt_pass -s 6 -c 6 -o /tmp/r02_prof_ntt $B > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
ncu -i /tmp/r02_prof_ntt.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r02_ncu_ntt_source.csv.gz
ls -la gpurun_out/r02_ncu_ntt_source.csv.gz
