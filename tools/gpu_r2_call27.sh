#!/bin/bash
# round 2, GPU call 27 (8 GPUs): concurrent host<->device copy rates with 1, 2, 4, 8 ranks active
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 tools/pcie_probe_ranks.py --gib 4 > gpurun_out/r02_pcie_ranks.txt 2> gpurun_out/pcie_ranks.err; echo "rc=$?"
cat gpurun_out/r02_pcie_ranks.txt; tail -3 gpurun_out/pcie_ranks.err
nvidia-smi topo -m 2>/dev/null | head -14 >> gpurun_out/r02_pcie_ranks.txt; lscpu | grep -E "Model name|Socket|NUMA node\(s\)|^CPU\(s\)" >> gpurun_out/r02_pcie_ranks.txt
