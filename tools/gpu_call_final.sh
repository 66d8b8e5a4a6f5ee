#!/bin/bash
# GPU call: parity tests, headline bench, data-independence sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep.py dist > gpurun_out/sweep_dist.jsonl 2> gpurun_out/sweep_dist.err; echo "sweep rc=$?"; cat gpurun_out/sweep_dist.jsonl; tail -3 gpurun_out/sweep_dist.err
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 800 gpurun_out/bench.err
cat gpurun_out/bench.json | cut -c1-600
