#!/bin/bash
# cfg5 (65536 cells x 128 components x 8192 samples) sharded by cell over N GPUs, both record exchanges
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for X in peer nccl; do
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-render --config cfg5 --queries -1 --exchange $X > gpurun_out/bench_cfg5_n${N}_$X.json 2> gpurun_out/bench_cfg5_n${N}_$X.err; echo "cfg5 N=$N $X rc=$?"; tail -2 gpurun_out/bench_cfg5_n${N}_$X.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_cfg5_n${N}_$X.json') if l.startswith('{')][0])
print('$X', d['value'], d['ms_per_step'], d['config'].get('exchange'), d['config'].get('peer_timeouts'))
PY
done
