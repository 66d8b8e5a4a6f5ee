#!/bin/bash
# GPU call: parity tests, headline bench (full line), data-independence sweep, then the ncu launch list of the bench command
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 400 gpurun_out/bench.err
cut -c1-700 gpurun_out/bench.json
timeout 600 python tools/sweep.py dist > gpurun_out/sweep_dist.jsonl 2> gpurun_out/sweep_dist.err; echo "sweep rc=$?"; cat gpurun_out/sweep_dist.jsonl
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_final2.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
