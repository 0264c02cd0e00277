set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'em_mstep' -s 8 -c 2 -f -o gpurun_out/prof_r2_mstep python tools/mstep_time.py > gpurun_out/ncu_mstep.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_mstep.log | cut -c1-150
