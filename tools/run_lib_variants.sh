#!/bin/bash
# Experiments: times the commit stages with each variants/lib_*.so (built by hand with extra -D flags) in place of the
# product library.  Run on the GPU box copy only (it overwrites xlayer_prover_b200/libglprover.so there).
for f in variants/lib_*.so; do
  cp $f xlayer_prover_b200/libglprover.so
  python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-verify 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; print('$f', 'commit %.1f intt %.1f ntt %.1f leaf %.1f' % (d['ms_per_step'], s['intt'], s['ntt_ext'], s['leaf_hash']))"
done
