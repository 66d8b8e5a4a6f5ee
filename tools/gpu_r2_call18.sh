#!/bin/bash
# round 2, GPU call 18: N = 1 bench on the final binary with the graded streamed commit (e2e and e2e_resident legs), full GPU suite
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench.err
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms'): print(k, d.get(k))
print('roofline', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'], 'resident', d.get('e2e_resident'))
"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400
