#!/usr/bin/env python
"""Where does the per-step covariance difference between the GPU and the reference come from?
For each stepwise iteration, from the REFERENCE's state: (a) the reference on the cell's samples in a permuted
order (pure float32 summation-order noise of the reference itself), (b) the GPU.  Prints the worst covariance
entry of each with the component's weight and mean shift.  usage: python tools/diag_step_noise.py [K N C iters]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import refbind  # noqa: E402
from parity import joint_param_errs  # noqa: E402
from sdmm_mitsuba_b200 import synth  # noqa: E402
from sdmm_mitsuba_b200.mixture import SDMMCells  # noqa: E402


def worst(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    live = b[:, 0, :] > 0
    cov = np.abs(a[:, 7:32] - b[:, 7:32]) / np.abs(b[:, 7:32]).max(axis=1, keepdims=True)
    cov = np.where(live[:, None, :], cov, 0)
    c, e, k = np.unravel_index(np.argmax(cov), cov.shape)
    return cov.max(), (int(c), int(e), int(k)), np.sort(cov.max(axis=1).ravel())[-5:]


def main():
    K, N, C, iters = (int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (64, 4096, 16, 4)))
    seed = 100 + K
    point6, weight, off = synth.make_samples(C, N, seed=seed)
    joint, em, scal = synth.make_initial_model(C, K, seed=seed + 1)
    rng = np.random.default_rng(0)
    perm = np.concatenate([c * N + rng.permutation(N) for c in range(C)])
    p6s, ws = np.ascontiguousarray(point6[:, perm]), np.ascontiguousarray(weight[perm])
    a = refbind.RefCells(C, K)
    a.set(joint, em, scal)
    b = refbind.RefCells(C, K)
    cells = SDMMCells(C, K)
    cells.upload(joint, em, scal)
    cells.set_samples(point6, weight, off)
    for it in range(iters):
        j0, e0, s0 = a.get()
        b.set(j0, e0, s0)
        cells.upload(j0, e0, s0)
        a.em_step(off, point6, weight, 1)
        b.em_step(off, p6s, ws, 1)
        cells.em_step(1, with_conditioner=False)
        j1 = a.get()[0]
        jb = b.get()[0]
        jg = cells.download()[0]
        for name, j in (("ref(permuted)", jb), ("gpu", jg)):
            m, (c, e, k), top = worst(j, j1)
            shift = np.abs(j1[c, 1:4, k] - j0[c, 1:4, k]).max()
            print("it %d %-14s cov err %.3g at cell %d comp %d entry %d | pmf %.3g, |cov|max %.3g, mean shift %.3g | top5 comps %s | %s"
                  % (it, name, m, c, k, e, j1[c, 0, k], np.abs(j1[c, 7:32, k]).max(), shift,
                     ["%.2g" % v for v in top], {q: float("%.3g" % v) for q, v in joint_param_errs(j, j1).items()}))


if __name__ == "__main__":
    main()
