#!/bin/bash
# round 2, GPU call 2: leaf-hash fold variants (form 2, wide MDS fold, pre-matrix unroll), parity tests with the new default
mkdir -p gpurun_out
( bash tools/ubench/run_variants.sh r1 vall v3s v3sm v4u1 v4u2 v4 v4u4 v4_mb4 v4_mb6 v4_b256 v4_b64 r1 v4 ) > gpurun_out/r02_pos_variants2.txt 2>&1
cat gpurun_out/r02_pos_variants2.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
