#!/bin/bash
set -u
mkdir -p gpurun_out
SWEEP_VARIANTS=${SWEEP_VARIANTS:-0,6,7} timeout 900 python tools/estep_sweep.py 64 2048 4096 2>&1 | tee gpurun_out/sweep_k64.log
