#!/bin/bash
# compute-sanitizer over K1 / K2+K3 / K4 (golden fixtures, small) and the two-context peer exchange.
# One tool per gpurun call (B200_PROFILING.md):  bash tools/gpu_sanitize.sh memcheck|racecheck
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
T="tests/test_gpu_golden.py::test_gpu_em_matches_golden tests/test_gpu_golden.py::test_gpu_queries_match_golden tests/test_gpu_peer.py::test_peer_push_matches_single_context tests/test_gpu_peer.py::test_peer_epochs_stay_in_step_when_ranks_pick_different_schedules"
timeout 600 python -m pytest $T -q -x -k "not n4096 and not n8192" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 2400 compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitizer_$TOOL.log python -m pytest $T -q -x -k "not n4096 and not n8192" > gpurun_out/sanitize_$TOOL.out 2>&1
echo "sanitizer rc=$?"; tail -3 gpurun_out/sanitize_$TOOL.out; tail -8 gpurun_out/sanitizer_$TOOL.log
