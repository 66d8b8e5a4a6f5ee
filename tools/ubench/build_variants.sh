#!/bin/bash
# usage: build_variants.sh name "flags" [name "flags" ...]  -> posv/<name> (+ posv/<name>.log with registers/spills)
cd "$(dirname "$0")"
mkdir -p posv
B="nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I../../xlayer_prover_b200/csrc -Xptxas -v"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( $B $flags -o posv/$name pos_bench.cu 2>&1 | grep -A2 "linear_hash" | grep -E "Used|spill" > posv/$name.log ) &
done
wait
head -3 posv/*.log
