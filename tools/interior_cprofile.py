"""cProfile of optimize() in the iteration of the interior/doubling loop that grows the tree (64 spp), second repetition."""
import sys, cProfile, pstats
sys.path.insert(0, ".")
import torch
from sdmm_mitsuba_b200 import scenes
from sdmm_mitsuba_b200.integrator import SDMMIntegrator
quads, cam = scenes.interior(640, 360)
pr = cProfile.Profile()
for rep in range(2):
    g = SDMMIntegrator(quads, cam, K=16, max_depth=6, samples_per_iteration=8, seed=5, pool_capacity=96_000_000)
    spp = 8
    for it in range(4):
        st, new = g.render_pass(spp, collect=True)
        torch.cuda.synchronize()
        if rep == 1 and it == 3:
            pr.enable()
        info = g.optimize(new)
        torch.cuda.synchronize()
        pr.disable()
        g.iteration += 1; g.total_spp += spp; spp *= 2
    g.close()
pstats.Stats(pr).sort_stats("tottime").print_stats(32)
