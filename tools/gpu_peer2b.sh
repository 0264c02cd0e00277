#!/bin/bash
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/mgpu_check.py 2>&1 | grep -v Warning | tail -3 | tee gpurun_out/mgpu_check_$N.log
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/bench_n${N}.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_n${N}.json') if l.startswith('{')][0])
print('N=$N', d['value'], d['ms_per_step'], d['config'].get('exchange'), d['config'].get('peer_timeouts'), d['gpu_launches'], d['e2e']['value'])
PY
head -c 60 gpurun_out/bench_n${N}.json; echo
