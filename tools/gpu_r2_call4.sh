#!/bin/bash
# round 2, GPU call 4 (2 GPUs): multi-device tests (glp_multi on distinct devices, sharded.py under torchrun, GLP_DEVICES drop-in), bench at N = 2
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -x -q -k "two_gpus or multi or dropin or group_proofs" > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu2.log
tail -15 gpurun_out/pytest_gpu2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"
tail -c 800 gpurun_out/bench_n2.err
python -c "
import json
d=json.load(open('gpurun_out/bench_n2.json'))
for k in ('value','ms_per_step','stage_ms','ntt_retries','verify','e2e'): print(k, d.get(k))
"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc=$?"
cut -c1-400 gpurun_out/bench_ref_n2.json
