#!/bin/bash
# GPU call: host topology next to PCIe copy rates with the process bound to each NUMA node
mkdir -p gpurun_out
{
nvidia-smi topo -m
lscpu | grep -iE "numa|socket|model name|^cpu\(s\)"
for d in /sys/bus/pci/devices/*; do if [ -f $d/class ] && grep -q "^0x0302\|^0x0300" $d/class 2>/dev/null; then echo "$d numa_node=$(cat $d/numa_node) local_cpulist=$(cat $d/local_cpulist)"; fi; done
cat /sys/devices/system/node/node*/cpulist 2>/dev/null
free -g | head -2
which numactl taskset
} > gpurun_out/numa_topo.txt 2>&1
cat gpurun_out/numa_topo.txt
cd tools/ubench && nvcc -O2 -o pcie_probe pcie_probe.cu 2>/dev/null; cd ../..
echo "=== default affinity" | tee gpurun_out/numa_pcie.txt
timeout 300 tools/ubench/pcie_probe 2>&1 | head -6 | tee -a gpurun_out/numa_pcie.txt
for n in /sys/devices/system/node/node*; do
  l=$(cat $n/cpulist); echo "=== taskset $l ($n)" | tee -a gpurun_out/numa_pcie.txt
  timeout 300 taskset -c $l tools/ubench/pcie_probe 2>&1 | head -6 | tee -a gpurun_out/numa_pcie.txt
done
