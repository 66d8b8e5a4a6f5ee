#!/bin/bash
# round 2, GPU call 5c: front-padded chunks (no tail copy), POS_PR_CHUNK 3..6
mkdir -p gpurun_out
( bash tools/ubench/run_variants.sh c1 c4 n3 n4 n5 n6 n4 ) > gpurun_out/r02_pos_variants3c.txt 2>&1
grep -E "^[cn][0-9]" gpurun_out/r02_pos_variants3c.txt
