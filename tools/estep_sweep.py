#!/usr/bin/env python
"""E-step shape sweep on one GPU: times sdmmb_em_step's E-step launches (CUDA events inside the
library) for the kernel variant selected by SDMMB_ESTEP_VARIANT.  One process per variant because the
library latches the variant at first launch.  usage: python tools/estep_sweep.py [K] [cells] [N]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(K, C, N):
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    from benchlegs.data import make_device_samples
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    dev = torch.device("cuda", 0)
    cells = SDMMCells(C, K)
    joint, em, scal = synth.make_initial_model(C, K, seed=1)
    cells.upload(joint, em, scal)
    p6, w = make_device_samples(torch, C, N, 7, dev)
    off = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)
    cells.set_samples(p6, w, off)
    for _ in range(2):
        cells.em_step(4)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        cells.em_step(4)
        ms, n = cells.last_estep_ms()
        best = min(best, ms / n)
    j, _, s = cells.download()
    print(json.dumps(dict(variant=os.environ.get("SDMMB_ESTEP_VARIANT", "0"), K=K, cells=C, N=N, estep_ms=best,
                          gupdates_per_s=C * K * N / best / 1e6, tflops=102.0 * C * K * N / best / 1e9,
                          finite=bool(np.isfinite(j).all()), pmf_sum=float(j[:, 0].sum(axis=1).mean()))), flush=True)


if __name__ == "__main__":
    if os.environ.get("_SWEEP_CHILD"):
        one(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))
    else:
        K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
        C = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
        N = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
        variants = os.environ.get("SWEEP_VARIANTS", "0,1,2,3,4,5").split(",")
        for v in variants:
            env = dict(os.environ, SDMMB_ESTEP_VARIANT=v, _SWEEP_CHILD="1")
            subprocess.call([sys.executable, os.path.abspath(__file__), str(K), str(C), str(N)], env=env)
