"""Per-iteration phase times of the interior/doubling loop (where does the host-side time go?)."""
import os, sys, time
sys.path.insert(0, ".")
import torch
from sdmm_mitsuba_b200 import scenes
from sdmm_mitsuba_b200.integrator import SDMMIntegrator
quads, cam = scenes.interior(640, 360)
REPS = int(os.environ.get('REPS', '2'))
for rep in range(REPS):
    g = SDMMIntegrator(quads, cam, K=16, max_depth=6, samples_per_iteration=8, seed=5, pool_capacity=96_000_000)
    spp, rendered = 8, 0
    while rendered < 504:
        torch.cuda.synchronize(); t0 = time.perf_counter()
        st, new = g.render_pass(spp, collect=True)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        info = g.optimize(new)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        if rep == REPS - 1:
            print("spp %3d render %.3f s optimize %.3f s pool %d overflow %d %s" % (spp, t1 - t0, t2 - t1, g.pool_count, int(st.pool_overflow), info))
        g.iteration += 1; g.total_spp += spp; rendered += spp; spp *= 2
    g.close()
