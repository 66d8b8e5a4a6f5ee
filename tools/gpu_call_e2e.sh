#!/bin/bash
# GPU call: where does the streamed (host-pointer) commit lose time against the device-resident one?
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv,noheader -lms 50 > gpurun_out/e2e_clk.txt &
SMI=$!
for m in full root tree; do
  echo "=== mode $m" | tee -a gpurun_out/e2e_probe.log
  E2E_MODE=$m GLP_TRACE=1 timeout 600 python tools/e2e_probe.py 23 660 96 >> gpurun_out/e2e_probe.log 2>&1
done
kill $SMI
sort gpurun_out/e2e_clk.txt | uniq -c | sort -rn | head -12
tail -60 gpurun_out/e2e_probe.log
