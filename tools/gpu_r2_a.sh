#!/bin/bash
# round 2, call A: parity of the un-centred E-step, operand-form micro-benchmarks, shape sweep, bench
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 120 tools/bin/ubench > gpurun_out/ubench_r2.txt 2>&1; echo "ubench rc=$?"; grep "warps/SMSP=4" gpurun_out/ubench_r2.txt | tail -12
timeout 900 python -m pytest tests/test_gpu_em.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest_em.log 2>&1; echo "pytest em rc=$?"; tail -15 gpurun_out/pytest_em.log
SWEEP_VARIANTS=0,1,2,3,4,5 timeout 600 python tools/estep_sweep.py 64 2048 4096 2>&1 | grep '^{' | tee gpurun_out/sweep_r2a.jsonl
SWEEP_VARIANTS=0,1 timeout 600 python tools/estep_sweep.py 128 1024 8192 2>&1 | grep '^{' | tee -a gpurun_out/sweep_r2a.jsonl
SWEEP_VARIANTS=0 timeout 600 python tools/estep_sweep.py 16 8192 4096 2>&1 | grep '^{' | tee -a gpurun_out/sweep_r2a.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','clocks','gpu_launches']})
print('e2e',d['e2e']['value'],d['e2e']['ms_per_step'])
print('roofline',d['roofline'])
PY
