#!/bin/bash
# round 2 ncu captures (one gpurun call = one profiler): launch list of the default bench path, then --set full of the
# E-step (K1), the M-step (K2+K3) and the query kernel (K4, binned and random order).  Each only after the same
# command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
P="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 --no-queries"
timeout 600 $P > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'em_|pack_|peer_' -c 200 \
    --csv --log-file gpurun_out/r2_launches_cfg2.csv $P > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 --no-queries --cells 1024"
timeout 600 $Q > gpurun_out/plain_small.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_estep' -s 13 -c 2 \
    -f -o gpurun_out/prof_r2_estep $Q > gpurun_out/ncu_estep.log 2>&1
echo "ncu estep rc=$?"; tail -2 gpurun_out/ncu_estep.log | cut -c1-160
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_mstep' -s 13 -c 2 \
    -f -o gpurun_out/prof_r2_mstep $Q > gpurun_out/ncu_mstep.log 2>&1
echo "ncu mstep rc=$?"; tail -2 gpurun_out/ncu_mstep.log | cut -c1-160
R="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 --queries 4194304 --cells 1024"
timeout 600 $R > gpurun_out/plain_q.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'query_pair' -c 8 \
    -f -o gpurun_out/prof_r2_query $R > gpurun_out/ncu_query.log 2>&1
echo "ncu query rc=$?"; tail -2 gpurun_out/ncu_query.log | cut -c1-160
ls -la gpurun_out/*.ncu-rep
