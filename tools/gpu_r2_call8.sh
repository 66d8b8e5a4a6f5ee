#!/bin/bash
# round 2, GPU call 8b: chunked leaf hash at the bench width (660 columns: steady state, warps spread over the whole loop) -- code size vs instruction count
mkdir -p gpurun_out
cd tools/ubench/posv
for v in c1 c3 c4 t3 t4 t4p1 t5 t6 c4_mb3; do timeout 120 ./$v 4194304 660 "$v w660"; done 2>&1 | tee ../../../gpurun_out/r02_pos_width2.txt
