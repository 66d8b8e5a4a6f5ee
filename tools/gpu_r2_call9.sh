#!/bin/bash
# round 2, GPU call 9: headline bench, then ncu: launch list of the bench command, --set full of the leaf hash and the NTT passes.
# The .ncu-rep files are turned into csv on the box and removed (gpurun_out/ is capped at 64 MiB).
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries','leaf_redo_rows'): print(k, d.get(k))
print('roofline', {k:d['roofline'][k] for k in ('achieved','peak','frac')}, 'sass', d['roofline_sass']['frac'], 'issue', d['roofline_issue']['frac'])
print('e2e', d['e2e']); print('resident', d.get('e2e_resident')); print(d['cpu_baseline']['value'], d['clocks'])
"
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_full.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_full.csv $B > gpurun_out/ncu_full.log 2>&1
echo "ncu list rc=$?"
$B --log-rows 18 > gpurun_out/plain_leaf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_linear_hash -s 1 -c 1 -o /tmp/r02_prof_leaf $B --log-rows 18 > gpurun_out/ncu_leaf.log 2>&1
echo "ncu leaf rc=$?"
$B --cols 64 > gpurun_out/plain_ntt.log 2>&1 && \
ncu --set full --clock-control none -k regex:k_ntt_pass -s 6 -c 6 -o /tmp/r02_prof_ntt $B --cols 64 > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
ncu -i /tmp/r02_prof_leaf.ncu-rep --page raw --csv > gpurun_out/r02_ncu_leaf_raw.csv 2>/dev/null
ncu -i /tmp/r02_prof_ntt.ncu-rep --page raw --csv > gpurun_out/r02_ncu_ntt_raw.csv 2>/dev/null
ncu -i /tmp/r02_prof_leaf.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r02_ncu_leaf_source.csv.gz
ls -la gpurun_out | tail -8; du -sh gpurun_out
