#!/bin/bash
# round 2, call B: K1 variant sweep (two-block phased variants) + full GPU test list without -x
set -u
mkdir -p gpurun_out
SWEEP_VARIANTS=0,6,7,8,9,10 timeout 600 python tools/estep_sweep.py 64 2048 4096 2>&1 | grep '^{' | tee gpurun_out/sweep_r2b.jsonl
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; grep -E "^FAILED|passed|failed" gpurun_out/pytest_gpu.log | tail -20
