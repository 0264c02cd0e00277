#!/bin/bash
mkdir -p gpurun_out
for FS in 1 2 3 4; do
SDMMB_PIPE_FIT_STREAMS=$FS timeout 600 python bench.py --no-cpu-baseline --no-render --queries -1 > gpurun_out/bench_fs$FS.json 2> gpurun_out/bench_fs$FS.err; echo "rc=$?"; tail -2 gpurun_out/bench_fs$FS.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_fs$FS.json'))
print('fit_streams=$FS', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
PY
done
timeout 600 python -m pytest tests/test_gpu_em.py -x -q -k "optimize_host" 2>&1 | tail -2
