#!/usr/bin/env python
"""Time of everything in sdmmb_em_step that is not the E-step kernel (M-steps + launch gaps), per iteration.
usage: python tools/mstep_time.py [K] [cells] [N]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from benchlegs.data import make_device_samples  # noqa: E402
from sdmm_mitsuba_b200 import synth  # noqa: E402
from sdmm_mitsuba_b200.mixture import SDMMCells  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
C = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
N = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
dev = torch.device("cuda", 0)
cells = SDMMCells(C, K)
joint, em, scal = synth.make_initial_model(C, K, seed=1)
cells.upload(joint, em, scal)
chunks = [make_device_samples(torch, min(1024, C - c0), N, 7 + c0, dev) for c0 in range(0, C, 1024)]
p6 = torch.cat([c[0] for c in chunks], dim=1).contiguous()
w = torch.cat([c[1] for c in chunks]).contiguous()
off = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)
cells.set_samples(p6, w, off)
for _ in range(3):
    cells.em_step(4)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
tot, est = [], []
for _ in range(10):
    e0.record()
    cells.em_step(4)
    e1.record()
    torch.cuda.synchronize()
    tot.append(e0.elapsed_time(e1))
    est.append(cells.last_estep_ms()[0])
print(json.dumps(dict(minb=os.environ.get("SDMMB_MSTEP_MINB", "2"), K=K, cells=C, N=N, step_ms=float(np.median(tot)),
                      estep_ms=float(np.median(est)), other_ms_per_iter=float(np.median(np.array(tot) - np.array(est))) / 4)))
