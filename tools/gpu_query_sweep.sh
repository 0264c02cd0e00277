#!/bin/bash
# K4 tile sweep at cfg3 (binned / random order)
for T in 512 1024 2048 4096; do
  SDMMB_QUERY_TILE=$T timeout 600 python bench.py --cells 256 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-render --no-cfg5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['queries']
print('tile $T binned %.3f ms random %.3f ms frac %.3f' % (d['binned']['ms'], d['random_order']['ms'], d['roofline']['frac']))"
done
