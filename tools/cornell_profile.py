"""Per-iteration phase times of the Cornell-box loop of bench.py (640x360, K=16, maxDepth 5, 64 spp per iteration, 512 spp,
training while spp < 1/4)."""
import os, sys, time
sys.path.insert(0, ".")
import torch
from sdmm_mitsuba_b200 import scenes
from sdmm_mitsuba_b200.integrator import SDMMIntegrator
quads, cam = scenes.cornell_box(640, 360)
REPS = int(os.environ.get("REPS", "2"))
for rep in range(REPS):
    g = SDMMIntegrator(quads, cam, K=16, max_depth=5, samples_per_iteration=64, seed=5, pool_capacity=8_000_000)
    rendered = 0
    while rendered < 512:
        still = rendered < 512 * 0.25
        torch.cuda.synchronize(); t0 = time.perf_counter()
        st, new = g.render_pass(64, collect=still)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        info = g.optimize(new) if still else {}
        torch.cuda.synchronize(); t2 = time.perf_counter()
        if rep == REPS - 1:
            print("it %d render %.4f s optimize %.4f s guided %d mis %d %s" % (g.iteration, t1 - t0, t2 - t1, st.guided_queries, st.mis_queries, info))
        g.iteration += 1; g.total_spp += 64; rendered += 64
    g.close()
