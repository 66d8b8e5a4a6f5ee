#!/bin/bash
# round 2, GPU call 19: remainder column group on the narrow mapping: parity, then LDE stage times with / without the split at the
# slice widths of 8 / 4 / 2 / 1 GPUs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "remainder or narrow or scatter or extend" > gpurun_out/pytest_split.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_split.log
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident"
: > gpurun_out/r02_split.txt
for c in 88 80 165 330 660; do
for ns in 1 0; do
GLP_NTT_NO_SPLIT=$ns timeout 600 $B --cols $c > gpurun_out/split_tmp.json 2> gpurun_out/split_tmp.err || { echo "bench cols $c nosplit $ns failed"; tail -3 gpurun_out/split_tmp.err; }
python -c "
import json
d=json.load(open('gpurun_out/split_tmp.json'))
s=d['stage_ms']
print('cols %4d  GLP_NTT_NO_SPLIT=%d  commit %8.2f ms  intt %7.2f  ntt_ext %7.2f  leaf %8.2f  launches %s' % ($c, $ns, d['ms_per_step'], s['intt'], s['ntt_ext'], s['leaf_hash'], d.get('gpu_launches')))
" | tee -a gpurun_out/r02_split.txt
done
done
