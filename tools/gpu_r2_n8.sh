#!/bin/bash
# round 2: the driver's scaling line at N = 8 (one node), then the reference arm under the same launcher
set -u
N=${1:-8}
mkdir -p gpurun_out
timeout 1700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench$N rc=$?"; tail -4 gpurun_out/bench_n$N.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.load(open('gpurun_out/bench_n%s.json'%N))
print({k:d[k] for k in ['value','ms_per_step','n_gpus','gpu_launches']}, d['config']['exchange'], d['config']['peer_timeouts'])
print('e2e',{k:d['e2e'][k] for k in ['value','ms_per_step','host_link']})
c=d['cfg5']; print('cfg5',{k:c[k] for k in c if k not in ('roofline','workload','samples')})
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>/dev/null | cut -c1-1200 | tee gpurun_out/bench_ref_n$N.json | cut -c1-300
nvidia-smi topo -m 2>/dev/null | head -14
