#!/bin/bash
# Round-end style check on one B200: GPU tests, smoke, the shim check, reference arm, own arm.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_gpu.log | tail -10
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
oracle/_ref/shim_check 64 48 2000 | tail -4
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 1500 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ['value','ms_per_step','clocks','gpu_launches']})
print('e2e',{k:d['e2e'][k] for k in ['value','ms_per_step','host_link','cells_with_nonzero_status']})
r=d['roofline']; print('roofline frac',r['frac'],'ms',r['ms_per_launch']); print(r.get('instruction_mix')); print(r.get('register_file_bound'))
print('parity',d['parity'])
print('cpu',d['cpu_baseline'])
c=d['cfg5']; print('cfg5',{k:c[k] for k in c if k not in ('roofline','workload','samples')}); print('cfg5 roofline frac',c['roofline']['frac'])
q=d['queries']; print('queries',q['binned'],q['random_order'],q['roofline']['frac'],q['e2e'])
print('render',{k:d['progressive_render'][k] for k in list(d['progressive_render'])[:8]})
PY
