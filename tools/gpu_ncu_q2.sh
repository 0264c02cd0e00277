#!/bin/bash
set -u
mkdir -p gpurun_out
Q="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --queries 16777216 --cells 4096"
timeout 600 $Q > gpurun_out/plain_q.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'query_|bin_' -c 40 \
    --csv --log-file gpurun_out/launches_q.csv $Q > gpurun_out/ncu_q.log 2>&1
echo "rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_q.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows: print(r[0], r[4][:60], r[-3], r[-1])
PY
