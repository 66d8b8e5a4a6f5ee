#!/bin/bash
# round 2, GPU call 12d: fold product on the ALU pipe instead of an IMAD.WIDE (rebalancing the pipes at +10 % instructions)
mkdir -p gpurun_out
cd tools/ubench/posv
for v in d5 f3; do timeout 120 ./$v 4194304 660 "$v w660"; done 2>&1 | tee ../../../gpurun_out/r02_pos_sync4.txt
