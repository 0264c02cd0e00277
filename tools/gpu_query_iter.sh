#!/bin/bash
# one iteration on K4: the query parity tests, then the bench line's queries block (cfg3: 2^24 vertices, K = 64)
set -o pipefail
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_query.py tests/test_gpu_golden.py tests/test_gpu_product.py tests/test_render.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --cells 256 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 > gpurun_out/q_iter.json 2> gpurun_out/q_iter.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/q_iter.json").read().strip().splitlines()[-1])
print(json.dumps(d.get("queries"), indent=None)[:1500])
PY
