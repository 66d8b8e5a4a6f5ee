#!/bin/bash
# round 2, GPU call 21: ncu source page (per-SASS-instruction execution counts) of the six NTT passes of one LDE at 2^23 x 64
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident --cols 64"
$B > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ntt_pass -s 6 -c 6 -o /tmp/r02_prof_ntt $B > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
ncu -i /tmp/r02_prof_ntt.ncu-rep --page raw --csv > gpurun_out/r02_ncu_ntt2_raw.csv 2>/dev/null
ncu -i /tmp/r02_prof_ntt.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip > gpurun_out/r02_ncu_ntt2_source.csv.gz
ls -la gpurun_out | tail -5; du -sh gpurun_out
