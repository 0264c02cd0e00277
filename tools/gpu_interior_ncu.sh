#!/bin/bash
# phase times of the interior/doubling loop, then ncu --set full of three late pt_bounce_kernel launches
set -u
mkdir -p gpurun_out
timeout 600 python tools/interior_profile.py > gpurun_out/interior_phases.log 2>&1; tail -7 gpurun_out/interior_phases.log
REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pt_bounce' -s ${SKIP:-40} -c 3 \
    -f -o gpurun_out/prof_r2_bounce python tools/interior_profile.py > gpurun_out/ncu_bounce.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bounce.log | cut -c1-160
