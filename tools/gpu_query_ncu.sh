#!/bin/bash
# ncu --set full of the query kernel (K4) alone; only after the same command exited 0 without ncu
set -u
mkdir -p gpurun_out
R="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 --queries 4194304 --cells 1024"
timeout 600 $R > gpurun_out/plain_q.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'query_|bin_' -c 14 \
    -f -o gpurun_out/prof_query_iter $R > gpurun_out/ncu_query.log 2>&1
echo "ncu query rc=$?"; tail -2 gpurun_out/ncu_query.log | cut -c1-160
