#!/bin/bash
# full default bench line (with the progressive-render leg) + ncu launch list of one short progressive loop
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_n1_full.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n1_full.log | cut -c1-300
R="python tools/render_check.py"
export W=320 H=180 SPI=32 SAMPLES=128 REF_SPP=64
timeout 300 $R > gpurun_out/render_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_render.csv $R > gpurun_out/ncu_render.log 2>&1
echo "ncu render rc=$?"
