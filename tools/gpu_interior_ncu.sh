#!/bin/bash
# ncu --set full of late launches of one tracer kernel in the interior/doubling loop: KERNEL=pt_finish SKIP=4 COUNT=2
set -u
mkdir -p gpurun_out
K=${KERNEL:-pt_bounce}
REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -s ${SKIP:-40} -c ${COUNT:-3} \
    -f -o gpurun_out/prof_r2_${TAG:-$K} python ${SCRIPT:-tools/interior_profile.py} > gpurun_out/ncu_$K.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$K.log | cut -c1-160
