#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','clocks','gpu_launches']})
print('e2e',d['e2e']['value'],d['e2e']['ms_per_step'])
print('roofline frac',d['roofline']['frac'],'traffic',d['roofline']['traffic'])
print('queries',json.dumps(d['queries'])[:1200])
PY
timeout 900 python bench.py --config cfg5 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --queries -1 > gpurun_out/bench_cfg5_n1.json 2> gpurun_out/bench_cfg5_n1.err; echo "cfg5 rc=$?"; tail -3 gpurun_out/bench_cfg5_n1.err; cut -c1-600 gpurun_out/bench_cfg5_n1.json
