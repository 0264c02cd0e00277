#!/bin/bash
# progressive-loop iteration: render parity tests + phase times of the interior/doubling loop
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_render.py tests/test_tree_levels.py -m gpu -x -q > gpurun_out/pytest_render.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_render.log
timeout 600 python tools/interior_profile.py > gpurun_out/interior_phases.log 2>&1; tail -7 gpurun_out/interior_phases.log
if [ "${NCU:-0}" = "1" ]; then
REPS=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/interior_launches.csv python tools/interior_profile.py > gpurun_out/interior_ncu.log 2>&1
python tools/launch_list_summary.py gpurun_out/interior_launches.csv | head -26
fi
