#!/bin/bash
# GPU call: parity tests, the widened-row sweep (quotient commit, FRI layers, constant tree), its ncu launch list and a
# --set full capture of the FRI fold
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep.py widened > gpurun_out/sweep_widened.jsonl 2> gpurun_out/sweep_widened.err; echo "sweep rc=$?"
cat gpurun_out/sweep_widened.jsonl; tail -5 gpurun_out/sweep_widened.err
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_widened.csv \
  python tools/sweep.py widened > gpurun_out/ncu_widened.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fri_fold -c 1 -o gpurun_out/prof_fri python tools/sweep.py widened > gpurun_out/ncu_fri.log 2>&1; echo "ncu fri rc=$?"
ls -la gpurun_out | tail -8
