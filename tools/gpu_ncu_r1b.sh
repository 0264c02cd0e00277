#!/bin/bash
# ncu captures of the default bench path (borrowed SoA binding): launch list of two steps + full-set capture
# of the E-step kernel.  Each capture only after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
P="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-render --queries -1"
timeout 600 $P > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'em_|pack_|peer_' -c 200 \
    --csv --log-file gpurun_out/launches_r1b.csv $P > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --queries -1 --cells 1024"
timeout 600 $Q > gpurun_out/plain_small.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_estep' -s 13 -c 2 \
    -f -o gpurun_out/prof_r1b $Q > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
