#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 600 python bench.py --no-cpu-baseline --no-render --queries -1 > gpurun_out/bench_e2e3.json 2> gpurun_out/bench_e2e3.err; echo "rc=$?"; tail -2 gpurun_out/bench_e2e3.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_e2e3.json'))
print('cached items', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['gpu_launches'])
PY
