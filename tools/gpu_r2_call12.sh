#!/bin/bash
# round 2, GPU call 12b: chunk size with the CTA barrier per permutation; 660- and 64-column rows
mkdir -p gpurun_out
cd tools/ubench/posv
for v in s3_128 s4_128 s5_128 s6_128 s4_64; do for w in 660 64 100; do timeout 120 ./$v 4194304 $w "$v w$w"; done; done 2>&1 | tee ../../../gpurun_out/r02_pos_sync2.txt
