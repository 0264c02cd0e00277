#!/bin/bash
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/mgpu_check.py 2>&1 | grep -v Warning | tail -3 | tee gpurun_out/mgpu_check_$N.log
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; tail -2 gpurun_out/bench_n$N.err
timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --config cfg5 --queries -1 > gpurun_out/bench_cfg5_n$N.json 2> gpurun_out/bench_cfg5_n$N.err; echo "bench cfg5 N=$N rc=$?"; tail -2 gpurun_out/bench_cfg5_n$N.err
cut -c1-330 gpurun_out/bench_n$N.json; echo; cut -c1-330 gpurun_out/bench_cfg5_n$N.json
