#!/bin/bash
# ncu captures: launch list of one full bench step, full-set capture of the E-step and of the query kernel
set -u
mkdir -p gpurun_out
P="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries 4194304"
timeout 600 $P > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'em_|pack_|query_' -c 200 \
    --csv --log-file gpurun_out/launches.csv $P > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries 2097152 --cells 1024"
timeout 600 $Q > gpurun_out/plain_small.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_estep|query_binned' -s 13 -c 4 \
    -f -o gpurun_out/prof_r1 $Q > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
