#!/bin/bash
# GPU call: overlapped-commit experiment (LDE of block k+1 under the leaf hash of block k) + Poseidon multiply/fold variants
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-verify"
run() { tag=$1; shift; echo "=== $tag" | tee -a gpurun_out/co.log; env "$@" timeout 300 $B 2>>gpurun_out/co.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if l.startswith('{'):
        d = json.loads(l); print(json.dumps({k: d[k] for k in ('ms_per_step', 'stage_ms', 'root', 'gpu_launches', 'clocks')}))
" | tee -a gpurun_out/co.log; }
run serial            GLP_CO=0
run co256x1           GLP_CO=1
run co256x2           GLP_CO=1 GLP_CO_CTAS=2
run co128x1           GLP_CO=1 GLP_CO_THREADS=128
run co128x2           GLP_CO=1 GLP_CO_THREADS=128 GLP_CO_CTAS=2
run co128x3           GLP_CO=1 GLP_CO_THREADS=128 GLP_CO_CTAS=3
run co512x1           GLP_CO=1 GLP_CO_THREADS=512
run full_prio         GLP_CO=2
run co256x1_c64       GLP_CO=1 GLP_CO_COLS=64
run co256x1_c160      GLP_CO=1 GLP_CO_COLS=160
run co256x2_c160      GLP_CO=1 GLP_CO_COLS=160 GLP_CO_CTAS=2
echo "=== poseidon variants" | tee -a gpurun_out/co.log
bash tools/ubench/run_variants.sh 2>&1 | tee -a gpurun_out/co.log
tail -5 gpurun_out/co.err
