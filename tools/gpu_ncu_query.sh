#!/bin/bash
set -u
mkdir -p gpurun_out
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries 2097152 --cells 1024"
timeout 600 $Q > gpurun_out/plain_small.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'query_pair|bin_' -c 8 \
    -f -o gpurun_out/prof_query $Q > gpurun_out/ncu_query.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_query.log
