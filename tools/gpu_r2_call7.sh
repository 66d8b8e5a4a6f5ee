#!/bin/bash
# round 2, GPU call 7: GPU tests, headline bench, then ncu: launch list of the bench command and --set full captures of the leaf hash and the NTT passes
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries','leaf_redo_rows'): print(k, d.get(k))
print('roofline', {k:d['roofline'][k] for k in ('achieved','peak','frac')}, 'sass', d['roofline_sass']['frac'], 'issue', d['roofline_issue']['frac'])
print('e2e', d['e2e']); print(d['measured_lane_ops_per_s']); print(d['cpu_baseline']['value'])
"
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_full.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_full.csv $B > gpurun_out/ncu_full.log 2>&1
echo "ncu list rc=$?"
$B --log-rows 18 > gpurun_out/plain_leaf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_linear_hash -s 1 -c 1 -o gpurun_out/r02_prof_leaf $B --log-rows 18 > gpurun_out/ncu_leaf.log 2>&1
echo "ncu leaf rc=$?"
$B --cols 64 > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ntt_pass -s 6 -c 6 -o gpurun_out/r02_prof_ntt $B --cols 64 > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
ls -la gpurun_out | tail -8
