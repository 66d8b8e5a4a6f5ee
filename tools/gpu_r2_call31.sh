#!/bin/bash
# round 2, GPU call 31: L2 prefetch of a later tile in the wide NTT passes: distance sweep (GLP_NTT_PREFETCH, quarters of the resident tile count)
mkdir -p gpurun_out
: > gpurun_out/r02_ntt_prefetch.txt
B="python bench.py --steps 4 --warmup 2 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident"
for pf in 0 2 4 8 16; do
GLP_NTT_PREFETCH=$pf timeout 600 $B > gpurun_out/pf_tmp.json 2> gpurun_out/pf_tmp.err || { echo "bench pf $pf failed"; tail -3 gpurun_out/pf_tmp.err; }
python -c "
import json
d=json.load(open('gpurun_out/pf_tmp.json'))
s=d['stage_ms']
print('GLP_NTT_PREFETCH=%-2d  commit %8.2f ms  intt %7.2f  ntt_ext %7.2f  leaf %8.2f' % ($pf, d['ms_per_step'], s['intt'], s['ntt_ext'], s['leaf_hash']))
" | tee -a gpurun_out/r02_ntt_prefetch.txt
done
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_random_shapes.py -m gpu -x -q -k "ntt or extend or remainder or scatter or lde or commit or random" > gpurun_out/pytest_ntt.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_ntt.log
