#!/bin/bash
# ncu launch list of one progressive loop (2 training iterations, then guided passes) and a full-set capture of
# the bounce kernel and of the query kernel inside it
set -u
mkdir -p gpurun_out
R="python tools/render_check.py"
export W=320 H=180 SPI=64 SAMPLES=384 REF_SPP=64
timeout 300 $R > gpurun_out/render_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
    --log-file gpurun_out/launches_render.csv $R > gpurun_out/ncu_render.log 2>&1
echo "ncu render list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pt_bounce|query_pair|init_cells' -s 40 -c 6 \
    -f -o gpurun_out/prof_render $R > gpurun_out/ncu_render_full.log 2>&1
echo "ncu render full rc=$?"; tail -2 gpurun_out/ncu_render_full.log
