#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_em.py tests/test_gpu_shim.py -x -q 2>&1 | tail -3
for CS in 2097152 1048576 4194304; do
SDMMB_PIPE_CHUNK_SAMPLES=$CS timeout 600 python bench.py --no-cpu-baseline --no-render --queries -1 > gpurun_out/bench_e2e_$CS.json 2> gpurun_out/bench_e2e_$CS.err; echo "rc=$?"; tail -2 gpurun_out/bench_e2e_$CS.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_e2e_$CS.json'))
print('chunk=$CS', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
PY
done
