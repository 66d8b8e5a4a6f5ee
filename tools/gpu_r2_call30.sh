#!/bin/bash
# round 2, GPU call 30 (2 GPUs): bench at N = 2 on the final binary, then the multi-device tests
mkdir -p gpurun_out
n=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench n$n rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_n$n.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries'): print(k, d.get(k))
print('verify', d['verify'].get('root_equals_n1'), 'e2e', d['e2e']['ms_per_step'])
"
timeout 900 python -m pytest tests -m gpu -x -q -k "two_gpus or dropin_same_bits or multi" > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu2.log
