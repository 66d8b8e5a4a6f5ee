#!/bin/bash
# builds the pos_bench variants listed below into ./posv/<name>
set -e
cd "$(dirname "$0")"
mkdir -p posv
build() { name=$1; shift; nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I../../xlayer_prover_b200/csrc "$@" -o posv/$name pos_bench.cu & }
build g4_roll_lock256   -DPOS_SBOX_GROUP=4  -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=1
build g4_roll_nolock256 -DPOS_SBOX_GROUP=4  -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=0
build g4_roll_lock512   -DPOS_SBOX_GROUP=4  -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=512 -DPOS_LEAF_LOCKSTEP=1
build g4_roll_lock128   -DPOS_SBOX_GROUP=4  -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=128 -DPOS_LEAF_LOCKSTEP=1
build g12_roll_lock256  -DPOS_SBOX_GROUP=12 -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=1
build g12_roll_nolock256 -DPOS_SBOX_GROUP=12 -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=0
build g6_roll_lock256   -DPOS_SBOX_GROUP=6  -DPOS_ROLL_PRE=1 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=1
build g12_unroll_lock256 -DPOS_SBOX_GROUP=12 -DPOS_ROLL_PRE=0 -DPOS_LEAF_BLOCK=256 -DPOS_LEAF_LOCKSTEP=1
build g12_unroll_nolock128 -DPOS_SBOX_GROUP=12 -DPOS_ROLL_PRE=0 -DPOS_LEAF_BLOCK=128 -DPOS_LEAF_LOCKSTEP=0
wait
ls -la posv
