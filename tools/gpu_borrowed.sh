#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_em.py -x -q 2>&1 | tail -3
for B in borrowed packed; do
timeout 600 python bench.py --no-cpu-baseline --no-render --no-e2e --queries -1 --binding $B > gpurun_out/bench_$B.json 2> gpurun_out/bench_$B.err; echo "rc=$?"; tail -2 gpurun_out/bench_$B.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_$B.json'))
print('$B', d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['frac'], d['roofline']['ms_per_launch'])
PY
done
