#!/bin/bash
# round 2, GPU call 25: NTT pass variants (CTA shape, unpredicated tile loads) on the GPU box copy of the tree
mkdir -p gpurun_out
cp xlayer_prover_b200/libglprover.so /tmp/lib_product.so
for f in variants/lib_*.so; do
  cp $f xlayer_prover_b200/libglprover.so
  python bench.py --steps 4 --warmup 2 --no-e2e --no-cpu-baseline --no-verify --no-e2e-resident 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms']; print('$f', 'commit %.1f intt %.2f ntt %.2f leaf %.1f' % (d['ms_per_step'], s['intt'], s['ntt_ext'], s['leaf_hash']))" | tee -a gpurun_out/r02_ntt_variants.txt
done
cp /tmp/lib_product.so xlayer_prover_b200/libglprover.so
