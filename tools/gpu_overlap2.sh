#!/bin/bash
mkdir -p gpurun_out
for M in 2 3 4; do
SDMMB_MSTEP_MINB=$M timeout 600 python bench.py --no-cpu-baseline --no-render --no-e2e --queries -1 > gpurun_out/bench_ovm$M.json 2> gpurun_out/bench_ovm$M.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_ovm$M.json'))
print('overlap minb=$M', d['value'], d['ms_per_step'], d['gpu_launches'])
PY
done
