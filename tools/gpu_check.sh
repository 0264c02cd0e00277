#!/bin/bash
# One gpurun call: GPU parity tests, smoke, default bench (both arms), ncu launch list + one full capture
# of the E-step kernel.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
P="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries 4194304"
timeout 600 $P > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'em_|pack_|query_' -c 200 \
    --csv --log-file gpurun_out/launches.csv $P > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries -1 --cells 1024"
timeout 600 $Q > gpurun_out/plain_small.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:em_estep -s 13 -c 2 \
    -f -o gpurun_out/prof_estep $Q > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench.json | cut -c1-1500
