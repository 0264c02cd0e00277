"""The reference's own EM micro-benchmark shape (sdmm/benchmarks/em.cpp:99-157: one cell, K = 64 identical
components, N = 10^6 identical samples on the mean, one EM step per benchmark iteration) on the GPU and -- the way
google-benchmark runs it -- on ONE host core with the reference's own code.  The reference records no result for
it; it is a timing cross-check on a degenerate input, not the headline."""
import time

import numpy as np


def run(torch, local, cpu=True, n_points=1_000_000, K=64, steps=10):
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    point6, weight, off, joint, em, scal = synth.make_reference_benchmark(n_points, K)
    dev = torch.device("cuda", local)
    cells = SDMMCells(1, K, device=local)
    cells.upload(joint, em, scal)
    p6 = torch.from_numpy(point6).to(dev)
    w = torch.from_numpy(weight).to(dev)
    cells.set_samples(p6, w, off)
    for _ in range(3):
        cells.em_step(1, with_conditioner=False)
    torch.cuda.synchronize(dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(steps):
        cells.em_step(1, with_conditioner=False)
    ev[1].record()
    torch.cuda.synchronize(dev)
    ms = ev[0].elapsed_time(ev[1]) / steps
    j, _, s = cells.download()
    cells.close()
    out = dict(workload="sdmm/benchmarks/em.cpp: 1 cell x %d identical components x %d identical samples, one EM step "
                        "per iteration (E-step + statistics + update_model + prepare)" % (K, n_points),
               ms_per_em_step=ms, value=float(K) * n_points / (ms * 1e-3), unit="updates/s",
               finite=bool(np.isfinite(j).all()), status_words_nonzero=int((s[:, 7] != 0).sum()),
               note="one cell = %d work items of <= 1024 samples: %.1f waves of the E-step's resident CTAs, a launch-bound "
                    "corner of the kernel" % ((n_points + 1023) // 1024, ((n_points + 1023) // 1024) / (148.0 * 3)))
    if cpu:
        try:
            from oracle import refbind
            if refbind.available():
                n_cpu = n_points // 4            # a bounded sample: the loop is linear in N
                o = refbind.RefCells(1, K)
                o.set(joint, em, scal)
                sub_off = np.array([0, n_cpu], np.uint32)
                p, ww = np.ascontiguousarray(point6[:, :n_cpu]), np.ascontiguousarray(weight[:n_cpu])
                o.em_batch(sub_off, p, ww, 1, with_prepare=False, n_threads=1)          # warm-up
                secs = o.em_batch(sub_off, p, ww, 2, with_prepare=False, n_threads=1) / 2
                out["cpu_baseline"] = dict(value=float(K) * n_cpu / secs, unit="updates/s", cores=1, kind="reference",
                                           ms_per_em_step_at_full_n=secs * 1e3 * (n_points / n_cpu),
                                           sample="the same input cut to %d samples, one core as the benchmark runs, %.2f s per step"
                                                  % (n_cpu, secs))
        except Exception as e:   # the leg is a cross-check: never take the bench line down with it
            out["cpu_baseline"] = dict(error=str(e))
    return out
