#!/bin/bash
# round 2, GPU call 3: full GPU test suite (new full-size tests, multi-context tests), headline bench in its new format, reference arm
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=12 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench.err
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries','leaf_redo_rows','cpu_baseline','verify'): print(k, d.get(k))
print('roofline', {k:d['roofline'][k] for k in ('achieved','peak','frac')}, 'sass', d['roofline_sass']['frac'], 'issue', d['roofline_issue']['frac'], 'e2e', d['e2e'])
print(d['measured_lane_ops_per_s'])
"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-1500 gpurun_out/bench_ref.json
nproc; free -g | head -2
