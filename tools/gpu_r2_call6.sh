#!/bin/bash
# round 2, GPU call 6: e2e (host buffers) timeline, download split over one / two copy streams, block widths
mkdir -p gpurun_out
( for split in 1 2; do for w in 96 160; do echo "=== GLP_D2H_SPLIT=$split width $w"; GLP_D2H_SPLIT=$split GLP_TRACE=1 python tools/e2e_probe.py 23 660 $w 2>&1 | tail -14; done; done ) > gpurun_out/r02_e2e_trace.txt 2>&1
grep -E "===|^width" gpurun_out/r02_e2e_trace.txt
