#!/bin/bash
# round 2, call C: full GPU test list, reference arm, the new bench line (N = 1)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_gpu.log | tail -20
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -25 gpurun_out/bench.err | grep -E "Elapsed|Maximum resident|Error|error|Traceback" ; tail -5 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','clocks','gpu_launches']})
print('e2e',{k:d['e2e'][k] for k in ['value','ms_per_step','host_link','cells_with_nonzero_status']})
r=d['roofline']; print('roofline frac',r['frac'],'ms',r['ms_per_launch']); print(r.get('instruction_mix')); print(r.get('register_file_bound'))
print('parity',d['parity'])
print('cpu',d['cpu_baseline'])
c=d['cfg5']; print('cfg5',{k:c[k] for k in c if k!='roofline'}); print('cfg5 roofline frac',c['roofline']['frac'],c['roofline'].get('register_file_bound'))
q=d['queries']; print('queries',q['binned'],q['random_order'],q['roofline']['frac'],q['e2e'])
print('status',d['config']['cells_with_nonzero_status'])
PY
