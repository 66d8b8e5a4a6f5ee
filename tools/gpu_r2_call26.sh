#!/bin/bash
# round 2, GPU call 26: full GPU suite + headline bench after the NTT CTA-shape change (8 warps x 3 CTAs, unpredicated loads)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench.err
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms'): print(k, d.get(k))
print('roofline', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'], 'resident', d.get('e2e_resident',{}).get('ms_per_step'), 'cpu', d['cpu_baseline']['value'])
"
python __graft_entry__.py smoke 2>&1 | tail -2
