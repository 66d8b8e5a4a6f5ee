#!/bin/bash
# round 2, GPU call 16: column-block plans of the streamed (host-pointer) commit at 2^23 x 660
mkdir -p gpurun_out
GLP_TRACE=1 timeout 1500 python tools/e2e_plan_probe.py 23 660 \
  "legacy:GLP_STREAM_GROW=0" \
  "graded150:" \
  "graded200:GLP_STREAM_GROW=200" \
  "mult32:GLP_STREAM_PLAN=32+64+96+160+160+152" \
  "fine:GLP_STREAM_PLAN=32+40+64+96+144+160+128" \
  "graded150w192:GLP_STREAM_COLS=192" \
  > gpurun_out/e2e_plan.log 2>&1; echo "rc=$?"
grep "best" gpurun_out/e2e_plan.log
