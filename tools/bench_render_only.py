"""The progressive-render leg of bench.py alone (no CPU baseline)."""
import json
import sys
sys.path.insert(0, ".")
import torch
from benchlegs import render_leg as bench_render
print(json.dumps(bench_render.run_render(torch, cpu=False)))
