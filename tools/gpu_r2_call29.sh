#!/bin/bash
# round 2, GPU call 29: guard-band test of the extension, then two more CTA shapes of the NTT passes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "remainder_column_group_split" > gpurun_out/pytest_guard.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_guard.log
bash tools/gpu_r2_call25.sh
