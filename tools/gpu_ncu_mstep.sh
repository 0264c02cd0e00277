#!/bin/bash
set -u
mkdir -p gpurun_out
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --queries -1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_mstep' -s 13 -c 2 \
    -f -o gpurun_out/prof_mstep $Q > gpurun_out/ncu_mstep.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_mstep.log | cut -c1-200
