#!/bin/bash
# round 2, GPU call 1: leaf-hash fold variants (micro-benchmark), parity tests, headline bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
( bash tools/ubench/run_variants.sh r1 vmin vwide vboth vboth_i2f vboth_pl vall r1 vall ) > gpurun_out/r02_pos_variants.txt 2>&1
cat gpurun_out/r02_pos_variants.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 400 gpurun_out/bench.err
cut -c1-900 gpurun_out/bench.json
