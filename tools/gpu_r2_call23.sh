#!/bin/bash
# round 2, GPU call 23: ncu source page of the leaf hash (2^19 rows x 660 columns) on the final binary
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident --log-rows 18"
$B > gpurun_out/plain_leaf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_linear_hash -s 1 -c 1 -o /tmp/r02_prof_leaf $B > gpurun_out/ncu_leaf.log 2>&1
echo "ncu leaf rc=$?"
ncu -i /tmp/r02_prof_leaf.ncu-rep --page source --csv --print-source sass 2>/dev/null | gzip > gpurun_out/r02_ncu_leaf_source.csv.gz
ls -la gpurun_out/r02_ncu_leaf_source.csv.gz
