#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
E2E_MODE=separate timeout 600 python tools/e2e_probe.py 23 660 96 2>&1 | tee gpurun_out/e2e_separate.log
