#!/bin/bash
# 2-GPU check of the peer-memory record exchange (CUDA IPC over NVLink) + bench with both exchanges
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 python -m pytest tests/test_gpu_peer.py -x -q 2>&1 | tail -3
timeout 300 $TR tools/mgpu_check.py 2>&1 | grep -v Warning | tail -5 | tee gpurun_out/mgpu_check_$N.log
for X in peer nccl; do
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-render --exchange $X > gpurun_out/bench_n${N}_$X.json 2> gpurun_out/bench_n${N}_$X.err; echo "bench N=$N $X rc=$?"; tail -3 gpurun_out/bench_n${N}_$X.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_n${N}_$X.json'))
print('$X', d['value'], d['ms_per_step'], d['config'].get('exchange'), d['config'].get('peer_timeouts'), d['gpu_launches'])
PY
done
