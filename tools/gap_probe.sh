#!/bin/bash
# commit time against the sum of its stages (idle gaps between kernels), with and without the nvidia-smi clock sampler
for ns in "" 1 "" 1 "" 1; do
  BENCH_NO_CLOCK_SAMPLER=$ns python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-verify 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; t=sum(v for k,v in s.items() if k!='note'); print('sampler off=[$ns] commit %.1f stages %.1f gap %.1f  (intt %.1f ntt %.1f leaf %.1f)' % (d['ms_per_step'], t, d['ms_per_step']-t, s['intt'], s['ntt_ext'], s['leaf_hash']))"
done
