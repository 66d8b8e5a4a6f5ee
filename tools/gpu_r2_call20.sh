#!/bin/bash
# round 2, GPU call 20: NTT passes with pointer-walking loads / stores / twiddle fetches: parity, then stage times
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_random_shapes.py -m gpu -x -q -k "ntt or extend or remainder or narrow or scatter or lde or commit or random or live" > gpurun_out/pytest_ntt.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_ntt.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident"
for c in 660 80; do
timeout 600 $B --cols $c > gpurun_out/ntt_tmp.json 2> gpurun_out/ntt_tmp.err || { echo "bench failed"; tail -3 gpurun_out/ntt_tmp.err; }
python -c "
import json
d=json.load(open('gpurun_out/ntt_tmp.json'))
s=d['stage_ms']
print('cols %4d  commit %8.2f ms  intt %7.2f  ntt_ext %7.2f  leaf %8.2f  retries %s' % ($c, d['ms_per_step'], s['intt'], s['ntt_ext'], s['leaf_hash'], d.get('ntt_retries')))
" | tee -a gpurun_out/r02_ntt_addr.txt
done
