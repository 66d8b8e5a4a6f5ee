#!/bin/bash
# round 2, GPU call 15: the allocation-free group-proof entry point alone in a fresh process, then the whole GPU suite (which
# ends with it), the headline bench on the final binary, and one ncu --set full of the leaf hash (csv made on the box).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/pytest_noalloc_alone.log
tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench.err
python -c "
import json
d=json.load(open('gpurun_out/bench.json'))
for k in ('value','ms_per_step','stage_ms'): print(k, d.get(k))
print('roofline', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'], 'resident', d.get('e2e_resident'))
"
B="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident"
$B --log-rows 18 > gpurun_out/plain_leaf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_linear_hash -s 1 -c 1 -o /tmp/r02_prof_leaf $B --log-rows 18 > gpurun_out/ncu_leaf.log 2>&1
echo "ncu leaf rc=$?"
ncu -i /tmp/r02_prof_leaf.ncu-rep --page raw --csv > gpurun_out/r02_ncu_leaf_final_raw.csv 2>/dev/null
python __graft_entry__.py smoke 2>&1 | tail -2
du -sh gpurun_out
