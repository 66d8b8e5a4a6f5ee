#!/bin/bash
# runs every binary in posv/ (or those named on the command line) on 2^22 x 64 and samples the SM clock meanwhile
cd "$(dirname "$0")/posv"
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv,noheader -lms 100 > /tmp/clk.txt &
SMI=$!
names="$@"; [ -z "$names" ] && names=$(ls | grep -v '\.log$')
for v in $names; do
  echo "--- $(tail -1 /tmp/clk.txt)"
  timeout 60 ./$v 4194304 64 $v
done
kill $SMI
echo "clock samples: $(sort /tmp/clk.txt | uniq -c | sort -rn | head -5)"
