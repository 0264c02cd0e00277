#!/bin/bash
# per-kernel durations of the random-order query path (binning, query, un-permute) at 4.19 M vertices
R="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 --queries 4194304 --cells 1024"
timeout 600 $R > gpurun_out/plain_q.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'query_|bin_' -c 40 --csv \
   --log-file gpurun_out/q_times.csv $R > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/q_times.csv")) if len(r) > 10]
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(list)
for r in rows[1:]:
    agg[(r[ix["Kernel Name"]][:60], r[ix["Metric Name"]])].append(float(r[ix["Metric Value"]].replace(",", "")))
for k, v in agg.items():
    print(k[0], k[1], "n=%d" % len(v), "last=%.1f" % v[-1], r[ix["Metric Unit"]] if False else "")
PY
