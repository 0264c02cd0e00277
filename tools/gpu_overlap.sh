#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_em.py tests/test_gpu_peer.py -x -q 2>&1 | tail -3
for O in "" "--no-overlap"; do
timeout 600 python bench.py --no-cpu-baseline --no-render --no-e2e --queries -1 $O > gpurun_out/bench_ov$O.json 2> gpurun_out/bench_ov$O.err; echo "rc=$?"; tail -2 gpurun_out/bench_ov$O.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_ov$O.json'))
print('overlap$O', d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['frac'], d['roofline']['ms_per_launch'], d['roofline']['share_of_step'])
PY
done
