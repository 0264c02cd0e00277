#!/bin/bash
# round 2, call D (2 GPUs): GPU tests on GPU 0, then the 2-rank bench line
set -u
mkdir -p gpurun_out
CUDA_VISIBLE_DEVICES=0 timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_gpu.log | tail -20
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_n2.json'))
print({k:d[k] for k in ['value','ms_per_step','n_gpus','gpu_launches']}, d['config']['exchange'], d['config']['peer_timeouts'], d['config']['schedule'][:40])
print('e2e',{k:d['e2e'][k] for k in ['value','ms_per_step','host_link']})
print('parity', d['parity']['stepwise'], d['parity']['free_running']['ll_err_p999'], d['parity']['free_running']['reference_order_noise'])
c=d['cfg5']; print('cfg5',{k:c[k] for k in c if k not in ('roofline','workload','samples')})
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | cut -c1-200
