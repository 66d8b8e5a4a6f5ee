#!/bin/bash
# GPU call (gpurun --gpus N): the headline bench at N GPUs, launched the way the driver launches it
N=${N:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.used --format=csv,noheader > gpurun_out/multi_gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 \
  > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?"
tail -c 1500 gpurun_out/bench_n$N.err
cat gpurun_out/bench_n$N.json | cut -c1-3000
