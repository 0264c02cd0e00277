#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; grep -E "^FAILED|^ERROR|passed|failed|^E  " gpurun_out/pytest_gpu.log | tail -20
oracle/_ref/shim_check 64 48 2000 | tail -6
