#!/bin/bash
# round 2, GPU call 17 (8 GPUs): final binary (chunked leaf hash, graded streamed commit): bench at N = 8 and 4
mkdir -p gpurun_out
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench n$n rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_n$n.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries','verify'): print(k, d.get(k))
print('e2e', d['e2e'])
"
done
